// Vocabulary-tree descent of ORBVocabulary::transform (reference Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1217-1259,
// called per feature from :1126-1194 by Frame::ComputeBoW, src/Frame.cc:521-528) and the all-pairs selection of
// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:241-306).  Both are FORB::distance / DescriptorDistance
// popcount work (Thirdparty/DBoW2/DBoW2/FORB.cpp:84-104 == src/ORBmatcher.cc:1728-1744).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace orbm {

// Device layout of the tree: "slots".  Slot 0 is the root; the children of every node occupy consecutive slots in
// the order of the reference's Node::children vector, so one descent step reads one contiguous run of 32-byte
// descriptors (k x 32 B, 320 B for ORBvoc) plus the {first child slot, child count} pair of each child, and the
// winner's pair is already in a register when the next level starts (one dependent load round per level).
struct VocabDev {
    const uint4* desc;        // [nslots][2]  node descriptor
    const uint2* kids;        // [nslots]     {first child slot, number of children}; 0 children = leaf (Node::isLeaf)
    const uint32_t* nodeId;   // [nslots]     reference node id (line order of the text file)
    const uint32_t* wordId;   // [nslots]     Node::word_id
    int L;                    // m_L
};

__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// Eight lanes per feature, four features per warp.  Lane c of a group scores children c, c+8, ...; the scan of
// :1236-1249 keeps the FIRST child with the strictly smallest distance = the minimum of (distance << 8 | child
// position) over the group.  out[i] = {word id, id of the ancestor at level L - levelsup (0 when that level <= 0 or
// the leaf is shallower)}.
constexpr int VOC_GROUP = 8;
__global__ void __launch_bounds__(256) vocab_descend_kernel(VocabDev v, const uint4* __restrict__ feat, size_t n, int levelsup,
                                                            uint2* __restrict__ out) {
    const int lane = threadIdx.x & 31, gl = lane & (VOC_GROUP - 1);
    const size_t f = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) / VOC_GROUP;
    const bool live = f < n;
    uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
    if (live) { a0 = feat[2 * f]; a1 = feat[2 * f + 1]; }
    const int nidLevel = v.L - levelsup;
    uint32_t slot = 0, nidSlot = 0;
    uint2 kd = live ? v.kids[0] : make_uint2(0, 0);
    int level = 0;
    while (__any_sync(0xffffffffu, kd.y != 0)) {
        level++;
        uint32_t best = 0xffffffffu;
        uint2 bestKids = make_uint2(0, 0);
        for (uint32_t c = gl; c < kd.y; c += VOC_GROUP) {
            const uint32_t s = kd.x + c;
            const uint4 b0 = v.desc[2 * (size_t)s], b1 = v.desc[2 * (size_t)s + 1];
            const uint2 ck = v.kids[s];
            const uint32_t key = ((uint32_t)hamming256(a0, a1, b0, b1) << 8) | c;      // c < 256 (m_k <= 20 in the text loader)
            if (key < best) { best = key; bestKids = ck; }
        }
#pragma unroll
        for (int s = VOC_GROUP / 2; s > 0; s >>= 1) {
            const uint32_t ob = __shfl_xor_sync(0xffffffffu, best, s);
            const uint32_t ox = __shfl_xor_sync(0xffffffffu, bestKids.x, s), oy = __shfl_xor_sync(0xffffffffu, bestKids.y, s);
            if (ob < best) { best = ob; bestKids = make_uint2(ox, oy); }
        }
        if (kd.y != 0) {                                     // groups that reached their leaf idle
            slot = kd.x + (best & 0xffu);
            kd = bestKids;
            if (level == nidLevel) nidSlot = slot;
        }
    }
    if (live && gl == 0) out[f] = make_uint2(v.wordId[slot], v.nodeId[nidSlot]);
}

// ---- MapPoint::ComputeDistinctiveDescriptors: one warp per map point.
// median of row i = element [0.5*(N-1)] of the sorted distances (:291-293) = the smallest value whose cumulative
// count exceeds k = (N-1)/2; the first row with the strictly smallest median wins (:295-299).
__global__ void __launch_bounds__(256) distinctive_kernel(const uint4* __restrict__ desc, const int32_t* __restrict__ start,
                                                          int npoints, int32_t* __restrict__ bestOut) {
    __shared__ uint32_t hist[8][288];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int p = blockIdx.x * 8 + w;
    if (p >= npoints) return;
    const int s0 = start[p], N = start[p + 1] - s0;
    if (N <= 0) { if (lane == 0) bestOut[p] = -1; return; }
    const int k = (N - 1) >> 1;                              // (int)(0.5*(N-1))
    const uint4* D = desc + 2 * (size_t)s0;
    int bestMedian = 0x7fffffff, bestIdx = 0;
    if (N <= 32) {
        uint4 m0 = make_uint4(0, 0, 0, 0), m1 = m0;
        if (lane < N) { m0 = D[2 * lane]; m1 = D[2 * lane + 1]; }
        for (int i = 0; i < N; i++) {
            uint4 r0, r1;
            r0.x = __shfl_sync(0xffffffffu, m0.x, i); r0.y = __shfl_sync(0xffffffffu, m0.y, i);
            r0.z = __shfl_sync(0xffffffffu, m0.z, i); r0.w = __shfl_sync(0xffffffffu, m0.w, i);
            r1.x = __shfl_sync(0xffffffffu, m1.x, i); r1.y = __shfl_sync(0xffffffffu, m1.y, i);
            r1.z = __shfl_sync(0xffffffffu, m1.z, i); r1.w = __shfl_sync(0xffffffffu, m1.w, i);
            const int d = lane < N ? hamming256(r0, r1, m0, m1) : 1 << 20;
            int lo = 0, hi = 256;                            // smallest v with #{d <= v} > k
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (__popc(__ballot_sync(0xffffffffu, d <= mid)) > k) hi = mid; else lo = mid + 1;
            }
            if (lo < bestMedian) { bestMedian = lo; bestIdx = i; }
        }
    } else {
        uint32_t* h = hist[w];
        for (int i = 0; i < N; i++) {
            for (int b = lane; b < 288; b += 32) h[b] = 0;
            __syncwarp();
            const uint4 r0 = D[2 * i], r1 = D[2 * i + 1];
            for (int j = lane; j < N; j += 32) atomicAdd(&h[hamming256(r0, r1, D[2 * j], D[2 * j + 1])], 1u);
            __syncwarp();
            uint32_t c[9], tot = 0;                          // lane owns bins 9*lane .. 9*lane+8
#pragma unroll
            for (int b = 0; b < 9; b++) { c[b] = h[9 * lane + b]; tot += c[b]; }
            uint32_t incl = tot;
#pragma unroll
            for (int s = 1; s < 32; s <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, incl, s); if (lane >= s) incl += o; }
            uint32_t cum = incl - tot;
            int med = 0x7fffffff;
#pragma unroll
            for (int b = 0; b < 9; b++) { cum += c[b]; if (med == 0x7fffffff && cum > (uint32_t)k) med = 9 * lane + b; }
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) med = min(med, __shfl_xor_sync(0xffffffffu, med, s));
            if (med < bestMedian) { bestMedian = med; bestIdx = i; }
            __syncwarp();
        }
    }
    if (lane == 0) bestOut[p] = bestIdx;
}

}  // namespace orbm
