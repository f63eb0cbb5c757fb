// Frame grid + windowed candidate search on the device:
//   Frame::PosInGrid / AssignFeaturesToGrid   reference src/Frame.cc:504-518, :292-311
//   Frame::GetFeaturesInArea                   src/Frame.cc:410-495
//   ORBmatcher::DescriptorDistance             src/ORBmatcher.cc:1728-1744
// For every query (window centre, radius, level range, 256-bit descriptor) the
// kernel enumerates the candidates of the target frame in the reference's order
// (grid column outer, grid row inner, insertion order inside a cell) and
// attaches the Hamming distance; the order-dependent acceptance logic of the
// Search* functions is replayed over these lists by the caller.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/orbm.h"

namespace orbm {

constexpr int GRID_CELLS = ORBM_GRID_COLS * ORBM_GRID_ROWS;

struct GridParams {
    float minX, minY, invW, invH;
};

struct WQuery {            // one window query
    float x, y, r;
    int32_t minLevel, maxLevel;
    int32_t pad[3];
    uint32_t desc[8];
};

// ---- grid build: one block per frame ---------------------------------------
// cellOf[i] = ix*48+iy or -1 (PosInGrid false); cellStart = exclusive scan of counts; items in ascending i
__global__ void __launch_bounds__(1024) grid_build_kernel(const orbx_keypoint* __restrict__ keysUn, int N, GridParams g,
                                                          int32_t* __restrict__ cellOf, int32_t* __restrict__ cellStart,
                                                          int32_t* __restrict__ items) {
    __shared__ int cnt[GRID_CELLS + 1];
    __shared__ int fill[GRID_CELLS];
    __shared__ int warpSum[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int c = tid; c <= GRID_CELLS; c += blockDim.x) cnt[c] = 0;
    __syncthreads();
    for (int i = tid; i < N; i += blockDim.x) {
        // posX = round((kp.pt.x-mnMinX)*mfGridElementWidthInv), :507-509 (float round, half away from zero)
        const int px = (int)roundf(__fmul_rn(__fsub_rn(keysUn[i].x, g.minX), g.invW));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(keysUn[i].y, g.minY), g.invH));
        int c = -1;
        if (!(px < 0 || px >= ORBM_GRID_COLS || py < 0 || py >= ORBM_GRID_ROWS)) { c = px * ORBM_GRID_ROWS + py; atomicAdd(&cnt[c], 1); }
        cellOf[i] = c;
    }
    __syncthreads();
    // exclusive scan of 3072 counts: 3 per thread at 1024 threads
    const int per = (GRID_CELLS + blockDim.x - 1) / blockDim.x;
    const int b = min(tid * per, GRID_CELLS), e = min(b + per, GRID_CELLS);
    int s = 0;
    for (int c = b; c < e; c++) s += cnt[c];
    int inc = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
    if (lane == 31) warpSum[wid] = inc;
    __syncthreads();
    int base = 0;
    for (int k = 0; k < wid; k++) base += warpSum[k];
    int run = base + inc - s;
    for (int c = b; c < e; c++) { const int v = cnt[c]; cnt[c] = run; cellStart[c] = run; run += v; }
    if (tid == blockDim.x - 1) { cnt[GRID_CELLS] = run; cellStart[GRID_CELLS] = run; }
    __syncthreads();
    // placement: scatter with a per-cell counter, then order each cell's few entries ascending = push_back order (:300-310)
    for (int c = tid; c < GRID_CELLS; c += blockDim.x) fill[c] = 0;
    __syncthreads();
    for (int i = tid; i < N; i += blockDim.x) {
        const int c = cellOf[i];
        if (c >= 0) items[cnt[c] + atomicAdd(&fill[c], 1)] = i;
    }
    __syncthreads();
    for (int c = tid; c < GRID_CELLS; c += blockDim.x) {
        const int n = fill[c];
        if (n < 2) continue;
        int32_t* it = items + cnt[c];
        for (int x = 1; x < n; x++) {                    // insertion sort of a handful of indices
            const int v = it[x];
            int y = x - 1;
            while (y >= 0 && it[y] > v) { it[y + 1] = it[y]; y--; }
            it[y + 1] = v;
        }
    }
}

struct CellRange { int x0, x1, y0, y1; bool empty; };

__device__ __forceinline__ CellRange window_cells(const WQuery& q, const GridParams& g) {
    // :430-447, float arithmetic in the reference's order
    CellRange c;
    c.empty = false;
    c.x0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(q.x, g.minX), q.r), g.invW)));
    if (c.x0 >= ORBM_GRID_COLS) c.empty = true;
    c.x1 = min(ORBM_GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(q.x, g.minX), q.r), g.invW)));
    if (c.x1 < 0) c.empty = true;
    c.y0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(q.y, g.minY), q.r), g.invH)));
    if (c.y0 >= ORBM_GRID_ROWS) c.empty = true;
    c.y1 = min(ORBM_GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(q.y, g.minY), q.r), g.invH)));
    if (c.y1 < 0) c.empty = true;
    return c;
}

// one warp per query.  FILL=false: counts[q] = number of candidates.  FILL=true: (index, distance) pairs at offs[q].
template <bool FILL>
__global__ void __launch_bounds__(256) window_kernel(const WQuery* __restrict__ queries, int nq, GridParams g,
                                                     const orbx_keypoint* __restrict__ keysUn,
                                                     const uint32_t* __restrict__ desc, const int32_t* __restrict__ cellStart,
                                                     const int32_t* __restrict__ items, int32_t* __restrict__ counts,
                                                     const int32_t* __restrict__ offs, int2* __restrict__ out, int outCap) {
    const int lane = threadIdx.x & 31;
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (qi >= nq) return;
    const WQuery q = queries[qi];
    const CellRange cr = window_cells(q, g);
    int n = 0;
    const int base = FILL ? offs[qi] : 0;
    if (!cr.empty) {
        const bool checkLevels = (q.minLevel > 0) || (q.maxLevel >= 0);      // :449
        for (int ix = cr.x0; ix <= cr.x1; ix++) {
            // cells (ix, y0..y1) are consecutive CSR rows: one contiguous item range, already in iy-then-insertion order
            const int s = cellStart[ix * ORBM_GRID_ROWS + cr.y0], e = cellStart[ix * ORBM_GRID_ROWS + cr.y1 + 1];
            for (int k0 = s; k0 < e; k0 += 32) {
                const int k = k0 + lane;
                bool ok = false;
                int idx = -1;
                if (k < e) {
                    idx = items[k];
                    const orbx_keypoint kp = keysUn[idx];
                    ok = true;
                    if (checkLevels) {
                        if (kp.octave < q.minLevel) ok = false;
                        if (q.maxLevel >= 0 && kp.octave > q.maxLevel) ok = false;
                    }
                    const float dx = __fsub_rn(kp.x, q.x), dy = __fsub_rn(kp.y, q.y);
                    if (!(fabsf(dx) < q.r && fabsf(dy) < q.r)) ok = false;      // square window, strict, :488
                }
                const unsigned m = __ballot_sync(0xffffffffu, ok);
                if (FILL && ok) {
                    const uint32_t* d = desc + (size_t)idx * 8;
                    int dist = 0;
#pragma unroll
                    for (int w = 0; w < 8; w++) dist += __popc(q.desc[w] ^ d[w]);
                    const int slot = base + n + __popc(m & ((1u << lane) - 1));
                    if (slot < outCap) out[slot] = make_int2(idx, dist);      // (the host re-runs with a larger buffer when offs[nq] > outCap)
                }
                n += __popc(m);
            }
        }
    }
    if (!FILL && lane == 0) counts[qi] = n;
}

// exclusive scan of counts (single block), total to offs[nq]
__global__ void __launch_bounds__(1024) scan_counts_kernel(const int32_t* __restrict__ counts, int n, int32_t* __restrict__ offs) {
    __shared__ int warpSum[32];
    __shared__ int carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int b0 = 0; b0 < n; b0 += 1024) {
        const int i = b0 + tid;
        const int v = i < n ? counts[i] : 0;
        int inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
        if (lane == 31) warpSum[wid] = inc;
        __syncthreads();
        int wb = 0;
        for (int k = 0; k < wid; k++) wb += warpSum[k];
        const int c0 = carry;
        if (i < n) offs[i] = c0 + wb + inc - v;
        __syncthreads();
        if (tid == 1023) carry = c0 + wb + inc;
        __syncthreads();
    }
    if (tid == 0) offs[n] = carry;
}

}  // namespace orbm
