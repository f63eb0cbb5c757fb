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
#include "pdl.cuh"

namespace orbx {

constexpr int CELL_MASK_ROWS = 72;  // CELL_MAX_DIM rounded up to whole 8-row batches
constexpr int CELL_MAX_DIM = 64;   // interior of a cell is < 60 px in each direction (wCell = ceil(W / floor(W/30)))

struct CellArgs {
    const uint8_t* score;             // frame 0 score block
    unsigned long long planeFrameStride;
    int nlevels;
    int iniTh, minTh;                 // the plane stores max(0, score + 1 - minTh)
    int totalCells;                   // per frame
    int cellBegin, cellEnd;           // cells this launch handles (a level range): [cellBegin, cellEnd)
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

// masks: this warp's shared scratch, [kind (0 = kept at minTh, 1 = kept at iniTh)][chunk][row] ballots
template <bool TWO>
__device__ __forceinline__ void cell_scan(const CellArgs& a, const uint8_t* __restrict__ S, int pitch, const CellBox b, int lane,
                                          uint32_t* __restrict__ out, int cap, uint32_t* __restrict__ cnt, int* err,
                                          unsigned (*masks)[2][CELL_MASK_ROWS]) {
    const int xa = b.x0 + lane, xb = b.x0 + 32 + lane;
    const bool ina = xa < b.x1, inb = TWO && xb < b.x1;
    const uint8_t* pa = S + ORB_XPAD + (size_t)b.y0 * pitch + xa;        // walks down the column
    const int rows = b.y1 - b.y0;
    const int iniTh = a.iniTh - a.minTh + 1;             // in the plane's encoding
    const uint32_t sBias = (uint32_t)(a.minTh - 1);
    // rolling state per chunk: centre value c, max(left,right) of its row lr, 3-max of the previous row hp
    int c0 = ina ? pa[0] : 0, c1 = inb ? pa[32] : 0;
    int lr0, lr1 = 0, hp0 = 0, hp1 = 0;
    {
        int l0, r0, l1, r1;
        if (TWO) { row_lr(c0, c1, lane, l0, r0, l1, r1); lr1 = max(l1, r1); }
        else { l0 = __shfl_up_sync(0xffffffffu, c0, 1); r0 = __shfl_down_sync(0xffffffffu, c0, 1); if (!lane) l0 = 0; if (lane == 31) r0 = 0; }
        lr0 = max(l0, r0);
    }
    constexpr int RB = TWO ? 4 : 8;                      // rows per batch; the next batch is fetched while this one is scanned
    int v0[RB], v1[RB];
    auto fetch = [&](int rb) {                           // rows rb+1 .. rb+RB (the "next" rows of steps rb .. rb+RB-1); zeros past the interior
#pragma unroll
        for (int i = 0; i < RB; i++) {
            const bool more = rb + i + 1 < rows;
            v0[i] = (ina && more) ? pa[(size_t)(i + 1) * pitch] : 0;
            if (TWO) v1[i] = (inb && more) ? pa[(size_t)(i + 1) * pitch + 32] : 0;
        }
        pa += (size_t)RB * pitch;
    };
    fetch(0);
    for (int rb = 0; rb < rows; rb += RB) {
        int n0v[RB], n1v[RB];
#pragma unroll
        for (int i = 0; i < RB; i++) { n0v[i] = v0[i]; if (TWO) n1v[i] = v1[i]; }
        if (rb + RB < rows) fetch(rb + RB);              // warp-uniform
#pragma unroll
        for (int i = 0; i < RB; i++) {
            const int r = rb + i;                        // rows >= `rows` see zeros only and leave empty masks (the mask rows exist)
            const int n0 = n0v[i], n1 = TWO ? n1v[i] : 0;
            int nl0, nr0, nl1 = 0, nr1 = 0;
            if (TWO) row_lr(n0, n1, lane, nl0, nr0, nl1, nr1);
            else { nl0 = __shfl_up_sync(0xffffffffu, n0, 1); nr0 = __shfl_down_sync(0xffffffffu, n0, 1); if (!lane) nl0 = 0; if (lane == 31) nr0 = 0; }
            // strictly greater than all 8 neighbours (non-corners and everything outside the interior count 0)
            const int nlr0 = max(nl0, nr0), hn0 = max(nlr0, n0);
            const bool k0 = c0 > max(max(hp0, hn0), lr0);
            const unsigned bm0 = __ballot_sync(0xffffffffu, k0), bi0 = __ballot_sync(0xffffffffu, k0 && c0 >= iniTh);
            if (lane == 0) { masks[0][0][r] = bm0; masks[1][0][r] = bi0; }
            hp0 = max(lr0, c0); c0 = n0; lr0 = nlr0;
            if (TWO) {
                const int nlr1 = max(nl1, nr1), hn1 = max(nlr1, n1);
                const bool k1 = c1 > max(max(hp1, hn1), lr1);
                const unsigned bm1 = __ballot_sync(0xffffffffu, k1), bi1 = __ballot_sync(0xffffffffu, k1 && c1 >= iniTh);
                if (lane == 0) { masks[0][1][r] = bm1; masks[1][1][r] = bi1; }
                hp1 = max(lr1, c1); c1 = n1; lr1 = nlr1;
            }
        }
    }
    __syncwarp();
    // totals at both thresholds: lane = row (rows <= 64: two rounds)
    unsigned mm[2][2], mi[2][2];                         // [round][chunk]
    int nMin = 0, nIni = 0;
#pragma unroll
    for (int q = 0; q < 2; q++) {
        const int r = lane + 32 * q;
        const bool in = r < rows;
        mm[q][0] = in ? masks[0][0][r] : 0u; mi[q][0] = in ? masks[1][0][r] : 0u;
        mm[q][1] = (TWO && in) ? masks[0][1][r] : 0u; mi[q][1] = (TWO && in) ? masks[1][1][r] : 0u;
        nMin += __popc(mm[q][0]) + __popc(mm[q][1]); nIni += __popc(mi[q][0]) + __popc(mi[q][1]);
    }
    int tMin = nMin, tIni = nIni;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { tMin += __shfl_xor_sync(0xffffffffu, tMin, d); tIni += __shfl_xor_sync(0xffffffffu, tIni, d); }
    const bool useIni = tIni > 0;                        // :976-987: iniThFAST first, minThFAST only if that found nothing
    const int total = useIni ? tIni : tMin;
    if (lane == 0) *cnt = (uint32_t)total;
    if (total == 0) return;
    if (total > cap) { if (lane == 0) atomicExch(err, 301); return; }
    // emission, pixel row-major: every lane writes the survivors of its own row(s) at the row's offset
    const uint8_t* col = S + ORB_XPAD;
    int carry = 0;
#pragma unroll
    for (int q = 0; q < 2; q++) {
        if (32 * q >= rows) break;                       // warp-uniform
        unsigned m0 = useIni ? mi[q][0] : mm[q][0], m1 = useIni ? mi[q][1] : mm[q][1];
        const int mine = __popc(m0) + __popc(m1);
        int incl = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        int pos = carry + incl - mine;
        carry += __shfl_sync(0xffffffffu, incl, 31);
        const int y = b.y0 + lane + 32 * q;
        const uint8_t* rowp = col + (size_t)y * pitch;
        while (m0) {
            const int x = b.x0 + __ffs(m0) - 1;
            m0 &= m0 - 1;
            out[pos++] = ((uint32_t)(x - ORB_BORDER_MIN)) | ((uint32_t)(y - ORB_BORDER_MIN) << 12) | (((uint32_t)rowp[x] + sBias) << 24);
        }
        if (TWO) while (m1) {
            const int x = b.x0 + 32 + __ffs(m1) - 1;
            m1 &= m1 - 1;
            out[pos++] = ((uint32_t)(x - ORB_BORDER_MIN)) | ((uint32_t)(y - ORB_BORDER_MIN) << 12) | (((uint32_t)rowp[x] + sBias) << 24);
        }
    }
}

__global__ void __launch_bounds__(256, 5) cell_nms_kernel(CellArgs a) {
    __shared__ unsigned s_masks[8][2][2][CELL_MASK_ROWS];
    const int lane = threadIdx.x & 31;
    const int cell = a.cellBegin + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int frame = blockIdx.y;
    if (cell >= a.cellEnd) return;
    unsigned (*masks)[2][CELL_MASK_ROWS] = s_masks[threadIdx.x >> 5];
    int l = 0;
    while (cell >= a.cellOff[l + 1]) l++;
    const int c = cell - a.cellOff[l];
    const int i = c / a.nCols[l], j = c - i * a.nCols[l];
    const CellBox b = cell_box(a, l, i, j);
    pdl_wait();                                          // the score planes are complete (few-frames schedule, pdl.cuh)
    pdl_launch_dependents();
    uint32_t* cnt = a.cellCount + (size_t)frame * a.totalCells + cell;
    if (!b.valid) { if (lane == 0) *cnt = 0; return; }
    if (b.x1 - b.x0 > CELL_MAX_DIM || b.y1 - b.y0 > CELL_MAX_DIM) { if (lane == 0) { *cnt = 0; atomicExch(a.err, 302); } return; }
    const uint8_t* S = a.score + (size_t)frame * a.planeFrameStride + a.planeOff[l];
    uint32_t* out = a.slots + (size_t)frame * a.slotFrameStride + a.slotOff[l] + (size_t)c * a.slotCap[l];
    if (b.x1 - b.x0 > 32) cell_scan<true>(a, S, a.pitch[l], b, lane, out, a.slotCap[l], cnt, a.err, masks);
    else cell_scan<false>(a, S, a.pitch[l], b, lane, out, a.slotCap[l], cnt, a.err, masks);
}

}  // namespace orbx
