// Per-cell corner selection of ComputeKeyPointsOctTree (reference
// src/ORBextractor.cc:952-1002): inside every ~30-px cell, 3x3 non-max
// suppression restricted to the cell's detection interior (cv::FAST is run on
// the cell sub-image, so scores outside the interior count as 0), keep the
// corners at iniThFAST, or - if the cell has none - those at minThFAST.
//
// One launch, one warp per cell: the score plane is read exactly once (lane =
// column, three rows rolling through registers, neighbours by warp shuffle),
// every row leaves two ballots (kept at minThFAST / kept at iniThFAST), and the
// cell's survivors are written in pixel row-major order into the cell's own
// slot list, packed x:12|y:12|score:8 with coordinates relative to (16,16).
// The quadtree kernel concatenates the slot lists in cell row-major order,
// which is exactly the order of vToDistributeKeys (:938-1002).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "geom.h"

namespace orbx {

constexpr int CELL_MAX_DIM = 64;   // interior of a cell is < 60 px in each direction (wCell = ceil(W / floor(W/30)))

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
    unsigned long long slotOff[ORB_MAX_LEVELS];   // element offset of the level's first cell slot list inside one frame
    int slotCap[ORB_MAX_LEVELS];                  // elements per cell slot list
    unsigned long long slotFrameStride;           // elements
    uint32_t* cellCount;              // [frames][totalCells]
    uint32_t* slots;                  // [frames][slotTotal]
    int* err;
};

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

// horizontal neighbours of a two-chunk row held as (c0, c1) per lane: chunk 0 covers columns x0..x0+31, chunk 1 the next 32
__device__ __forceinline__ void row_lr(int c0, int c1, int lane, int& l0, int& r0, int& l1, int& r1) {
    const int up0 = __shfl_up_sync(0xffffffffu, c0, 1), dn0 = __shfl_down_sync(0xffffffffu, c0, 1);
    const int up1 = __shfl_up_sync(0xffffffffu, c1, 1), dn1 = __shfl_down_sync(0xffffffffu, c1, 1);
    const int last0 = __shfl_sync(0xffffffffu, c0, 31), first1 = __shfl_sync(0xffffffffu, c1, 0);
    l0 = lane ? up0 : 0;
    r0 = lane < 31 ? dn0 : first1;
    l1 = lane ? up1 : last0;
    r1 = lane < 31 ? dn1 : 0;
}

template <bool TWO>
__device__ __forceinline__ void cell_scan(const CellArgs& a, const uint8_t* __restrict__ S, int pitch, const CellBox b, int lane,
                                          uint32_t* __restrict__ out, int cap, uint32_t* __restrict__ cnt, int* err) {
    // per-lane ballots of the rows lane and lane+32 (interior height < 64)
    unsigned mMin[2][2] = {{0, 0}, {0, 0}}, mIni[2][2] = {{0, 0}, {0, 0}};   // [chunk][row >> 5]
    int nIni = 0, nMin = 0;
    const int xa = b.x0 + lane, xb = b.x0 + 32 + lane;
    const bool ina = xa < b.x1, inb = TWO && xb < b.x1;
    const uint8_t* col = S + ORB_XPAD;
    auto ld = [&](int y, int& c0, int& c1) {
        c0 = (ina && y < b.y1) ? col[(size_t)y * pitch + xa] : 0;
        c1 = (inb && y < b.y1) ? col[(size_t)y * pitch + xb] : 0;
    };
    int p0 = 0, p1 = 0, hp0 = 0, hp1 = 0;             // previous row and its horizontal 3-max
    int c0, c1, n0, n1;
    ld(b.y0, c0, c1);
    int cl0, cr0, cl1 = 0, cr1 = 0;
    if (TWO) row_lr(c0, c1, lane, cl0, cr0, cl1, cr1);
    else { cl0 = __shfl_up_sync(0xffffffffu, c0, 1); cr0 = __shfl_down_sync(0xffffffffu, c0, 1); if (!lane) cl0 = 0; if (lane == 31) cr0 = 0; }
    const int rows = b.y1 - b.y0;
    for (int r = 0; r < rows; r++) {
        const int y = b.y0 + r;
        ld(y + 1, n0, n1);                              // zeros past the last interior row
        int nl0, nr0, nl1 = 0, nr1 = 0;
        if (TWO) row_lr(n0, n1, lane, nl0, nr0, nl1, nr1);
        else { nl0 = __shfl_up_sync(0xffffffffu, n0, 1); nr0 = __shfl_down_sync(0xffffffffu, n0, 1); if (!lane) nl0 = 0; if (lane == 31) nr0 = 0; }
        const int hc0 = max(max(cl0, cr0), c0), hn0 = max(max(nl0, nr0), n0);
        // strictly greater than all 8 neighbours (non-corners and everything outside the interior count 0)
        const bool k0 = c0 > max(max(hp0, hn0), max(cl0, cr0));
        const unsigned bm0 = __ballot_sync(0xffffffffu, k0), bi0 = __ballot_sync(0xffffffffu, k0 && c0 >= a.iniTh);
        nMin += __popc(bm0); nIni += __popc(bi0);
        if (lane == (r & 31)) { mMin[0][0] = (r < 32) ? bm0 : mMin[0][0]; mMin[0][1] = (r >= 32) ? bm0 : mMin[0][1];
                                mIni[0][0] = (r < 32) ? bi0 : mIni[0][0]; mIni[0][1] = (r >= 32) ? bi0 : mIni[0][1]; }
        hp0 = hc0; p0 = c0; c0 = n0; cl0 = nl0; cr0 = nr0;
        if (TWO) {
            const int hc1 = max(max(cl1, cr1), c1), hn1 = max(max(nl1, nr1), n1);
            const bool k1 = c1 > max(max(hp1, hn1), max(cl1, cr1));
            const unsigned bm1 = __ballot_sync(0xffffffffu, k1), bi1 = __ballot_sync(0xffffffffu, k1 && c1 >= a.iniTh);
            nMin += __popc(bm1); nIni += __popc(bi1);
            if (lane == (r & 31)) { mMin[1][0] = (r < 32) ? bm1 : mMin[1][0]; mMin[1][1] = (r >= 32) ? bm1 : mMin[1][1];
                                    mIni[1][0] = (r < 32) ? bi1 : mIni[1][0]; mIni[1][1] = (r >= 32) ? bi1 : mIni[1][1]; }
            hp1 = hc1; p1 = c1; c1 = n1; cl1 = nl1; cr1 = nr1;
        }
    }
    (void)p0; (void)p1;
    const bool useIni = nIni > 0;                       // :976-987: iniThFAST first, minThFAST only if that found nothing
    const int total = useIni ? nIni : nMin;
    if (lane == 0) *cnt = (uint32_t)total;
    if (total == 0) return;
    if (total > cap) { if (lane == 0) atomicExch(err, 301); return; }
    int pos = 0;
    for (int r = 0; r < rows; r++) {
        const int y = b.y0 + r;
        const unsigned w0 = useIni ? ((r < 32) ? mIni[0][0] : mIni[0][1]) : ((r < 32) ? mMin[0][0] : mMin[0][1]);
        const unsigned m0 = __shfl_sync(0xffffffffu, w0, r & 31);
        if (m0) {
            if ((m0 >> lane) & 1u) {
                const uint32_t s = col[(size_t)y * pitch + xa];
                out[pos + __popc(m0 & ((1u << lane) - 1))] =
                    ((uint32_t)(xa - ORB_BORDER_MIN)) | ((uint32_t)(y - ORB_BORDER_MIN) << 12) | (s << 24);
            }
            pos += __popc(m0);
        }
        if (TWO) {
            const unsigned w1 = useIni ? ((r < 32) ? mIni[1][0] : mIni[1][1]) : ((r < 32) ? mMin[1][0] : mMin[1][1]);
            const unsigned m1 = __shfl_sync(0xffffffffu, w1, r & 31);
            if (m1) {
                if ((m1 >> lane) & 1u) {
                    const uint32_t s = col[(size_t)y * pitch + xb];
                    out[pos + __popc(m1 & ((1u << lane) - 1))] =
                        ((uint32_t)(xb - ORB_BORDER_MIN)) | ((uint32_t)(y - ORB_BORDER_MIN) << 12) | (s << 24);
                }
                pos += __popc(m1);
            }
        }
    }
}

__global__ void __launch_bounds__(256) cell_nms_kernel(CellArgs a) {
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
    if (!b.valid) { if (lane == 0) *cnt = 0; return; }
    if (b.x1 - b.x0 > CELL_MAX_DIM || b.y1 - b.y0 > CELL_MAX_DIM) { if (lane == 0) { *cnt = 0; atomicExch(a.err, 302); } return; }
    const uint8_t* S = a.score + (size_t)frame * a.planeFrameStride + a.planeOff[l];
    uint32_t* out = a.slots + (size_t)frame * a.slotFrameStride + a.slotOff[l] + (size_t)c * a.slotCap[l];
    if (b.x1 - b.x0 > 32) cell_scan<true>(a, S, a.pitch[l], b, lane, out, a.slotCap[l], cnt, a.err);
    else cell_scan<false>(a, S, a.pitch[l], b, lane, out, a.slotCap[l], cnt, a.err);
}

}  // namespace orbx
