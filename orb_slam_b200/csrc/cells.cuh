// Per-cell corner selection of ComputeKeyPointsOctTree (reference
// src/ORBextractor.cc:952-1002): inside every ~30-px cell, 3x3 non-max
// suppression restricted to the cell's detection interior (cv::FAST is run on
// the cell sub-image, so scores outside the interior count as 0), keep the
// corners at iniThFAST, or - if the cell has none - those at minThFAST.
// Output: vToDistributeKeys in the reference's order (cell row-major, then
// pixel row-major), coordinates relative to (16,16), packed x:12|y:12|score:8.
// Three launches: count per cell -> exclusive scan per (frame, level) -> emit.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "geom.h"

namespace orbx {

struct CellArgs {
    const uint8_t* score;             // frame 0 score block
    unsigned long long planeFrameStride;
    int nlevels;
    int iniTh;
    int totalCells;                   // per frame
    // per level
    int cellOff[ORB_MAX_LEVELS + 1];
    int nCols[ORB_MAX_LEVELS], wCell[ORB_MAX_LEVELS], hCell[ORB_MAX_LEVELS];
    int maxBX[ORB_MAX_LEVELS], maxBY[ORB_MAX_LEVELS];
    int pitch[ORB_MAX_LEVELS];
    unsigned long long planeOff[ORB_MAX_LEVELS];
    unsigned long long candOff[ORB_MAX_LEVELS];
    int candCap[ORB_MAX_LEVELS];
    unsigned long long candFrameStride;   // elements
    uint32_t* cellCount;              // [frames][totalCells]: count | (useMin << 31); after the scan: offsets in cellOffs
    uint32_t* cellOffs;               // [frames][totalCells]
    uint32_t* cand;                   // [frames][candTotal]
    int* candCount;                   // [frames][nlevels]
};

constexpr int CELL_ROWS = 8;   // rows of a cell column loaded together

struct CellBox { int x0, x1, y0, y1; bool valid; };

// detection interior of cell (i,j) in level coordinates, :955-972 (+3/-3 of cv::FAST)
__device__ __forceinline__ CellBox cell_box(const CellArgs& a, int l, int i, int j) {
    CellBox b;
    const int iniY = ORB_BORDER_MIN + i * a.hCell[l];
    const int iniX = ORB_BORDER_MIN + j * a.wCell[l];
    int maxY = iniY + a.hCell[l] + 6, maxX = iniX + a.wCell[l] + 6;
    b.valid = !(iniY >= a.maxBY[l] - 3) && !(iniX >= a.maxBX[l] - 6);   // asymmetric on purpose (:958,:969)
    if (maxY > a.maxBY[l]) maxY = a.maxBY[l];
    if (maxX > a.maxBX[l]) maxX = a.maxBX[l];
    b.x0 = iniX + 3; b.x1 = maxX - 3; b.y0 = iniY + 3; b.y1 = maxY - 3;
    if (b.x1 <= b.x0 || b.y1 <= b.y0) b.valid = false;   // sub-image narrower than 7: cv::FAST finds nothing
    return b;
}

__device__ __forceinline__ bool nms_keep(const uint8_t* __restrict__ S, int pitch, int x, int y, int s, const CellBox& b) {
    // strictly greater than all 8 neighbours; neighbours outside the interior count 0 (s > 0 already)
    const bool l = x > b.x0, r = x + 1 < b.x1, u = y > b.y0, d = y + 1 < b.y1;
    const uint8_t* p = S + (size_t)y * pitch + ORB_XPAD + x;
    if (l && p[-1] >= s) return false;
    if (r && p[1] >= s) return false;
    if (u) {
        const uint8_t* q = p - pitch;
        if (q[0] >= s || (l && q[-1] >= s) || (r && q[1] >= s)) return false;
    }
    if (d) {
        const uint8_t* q = p + pitch;
        if (q[0] >= s || (l && q[-1] >= s) || (r && q[1] >= s)) return false;
    }
    return true;
}

// one warp per cell; EMIT=false counts, EMIT=true writes candidates at the scanned offsets
template <bool EMIT>
__global__ void __launch_bounds__(256) cell_kernel(CellArgs a) {
    const int lane = threadIdx.x & 31;
    const int cell = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int frame = blockIdx.y;
    if (cell >= a.totalCells) return;
    int l = 0;
    while (cell >= a.cellOff[l + 1]) l++;
    const int c = cell - a.cellOff[l];
    const int i = c / a.nCols[l], j = c - i * a.nCols[l];
    const CellBox b = cell_box(a, l, i, j);
    uint32_t* cnt = a.cellCount + (size_t)frame * a.totalCells + cell;
    if (!b.valid) { if (!EMIT && lane == 0) *cnt = 0; return; }
    const uint8_t* S = a.score + (size_t)frame * a.planeFrameStride + a.planeOff[l];
    const int pitch = a.pitch[l];
    if (!EMIT) {
        int nIni = 0, nMin = 0;
        for (int xb = b.x0; xb < b.x1; xb += 32) {
            const int x = xb + lane;
            for (int yb = b.y0; yb < b.y1; yb += CELL_ROWS) {
                int sv[CELL_ROWS];
#pragma unroll
                for (int r = 0; r < CELL_ROWS; r++)       // independent loads first: CELL_ROWS rows in flight
                    sv[r] = (x < b.x1 && yb + r < b.y1) ? S[(size_t)(yb + r) * pitch + ORB_XPAD + x] : 0;
#pragma unroll
                for (int r = 0; r < CELL_ROWS; r++) {
                    const int s = sv[r];
                    const bool keep = s > 0 && nms_keep(S, pitch, x, yb + r, s, b);
                    nMin += __popc(__ballot_sync(0xffffffffu, keep));
                    nIni += __popc(__ballot_sync(0xffffffffu, keep && s >= a.iniTh));
                }
            }
        }
        if (lane == 0) *cnt = nIni > 0 ? (uint32_t)nIni : ((uint32_t)nMin | 0x80000000u);
    } else {
        const uint32_t cv = *cnt;
        if ((cv & 0x7fffffffu) == 0) return;
        const int th = (cv & 0x80000000u) ? 1 : a.iniTh;      // the score plane is already 0 below minThFAST
        uint32_t pos = a.cellOffs[(size_t)frame * a.totalCells + cell];
        uint32_t* out = a.cand + (size_t)frame * a.candFrameStride + a.candOff[l];
        // cells are < 60 px wide (wCell = ceil(W / floor(W/30))): at most two 32-lane chunks per row
        const int nch = (b.x1 - b.x0 > 32) ? 2 : 1;
        for (int yb = b.y0; yb < b.y1; yb += CELL_ROWS) {
            int sv[CELL_ROWS][2];
#pragma unroll
            for (int r = 0; r < CELL_ROWS; r++)
#pragma unroll
                for (int c2 = 0; c2 < 2; c2++) {
                    const int x = b.x0 + 32 * c2 + lane;
                    sv[r][c2] = (c2 < nch && x < b.x1 && yb + r < b.y1) ? S[(size_t)(yb + r) * pitch + ORB_XPAD + x] : 0;
                }
#pragma unroll
            for (int r = 0; r < CELL_ROWS; r++)
#pragma unroll
                for (int c2 = 0; c2 < 2; c2++) {
                    if (c2 >= nch) continue;                 // warp-uniform
                    const int x = b.x0 + 32 * c2 + lane, y = yb + r;
                    const int s = sv[r][c2];
                    const bool keep = s >= th && nms_keep(S, pitch, x, y, s, b);
                    const unsigned m = __ballot_sync(0xffffffffu, keep);
                    if (keep) {
                        const uint32_t k = pos + __popc(m & ((1u << lane) - 1));
                        out[k] = ((uint32_t)(x - ORB_BORDER_MIN)) | ((uint32_t)(y - ORB_BORDER_MIN) << 12) | ((uint32_t)s << 24);
                    }
                    pos += __popc(m);
                }
        }
    }
}

// exclusive scan of the cell counts of one (frame, level); block per (level, frame)
__global__ void __launch_bounds__(256) cell_scan_kernel(CellArgs a) {
    __shared__ uint32_t warpSum[8];
    __shared__ uint32_t carry;
    const int l = blockIdx.x, frame = blockIdx.y;
    const int n = a.cellOff[l + 1] - a.cellOff[l];
    const uint32_t* cnt = a.cellCount + (size_t)frame * a.totalCells + a.cellOff[l];
    uint32_t* offs = a.cellOffs + (size_t)frame * a.totalCells + a.cellOff[l];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 256) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < n ? (cnt[i] & 0x7fffffffu) : 0;
        uint32_t inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += t;
        }
        if (lane == 31) warpSum[wid] = inc;
        __syncthreads();
        uint32_t wbase = 0;
        for (int k = 0; k < wid; k++) wbase += warpSum[k];
        const uint32_t c0 = carry;
        if (i < n) offs[i] = c0 + wbase + inc - v;
        __syncthreads();
        if (threadIdx.x == 255) carry = c0 + wbase + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) a.candCount[frame * a.nlevels + l] = (int)carry;
}

}  // namespace orbx
