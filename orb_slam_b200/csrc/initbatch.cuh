// ORBmatcher::SearchForInitialization (reference src/ORBmatcher.cc:417-588) for MANY frame pairs at once, entirely on the
// device: the frames are the extractor's batch output as it lies in HBM (cv::KeyPoint[frames][cap], descriptors
// [frames][cap][32], counts[frames]); nothing travels to the host.  BASELINE.json configs[2] (EuRoC batch, frame i against
// frame i+1, window 100, nnratio 0.9, rotation histogram).
//   grid_build_batch_kernel : Frame::AssignFeaturesToGrid / PosInGrid (src/Frame.cc:292-311, :504-518), one block per frame
//   init_match_batch_kernel : one warp per pair replays the function's loop in query order - the queries of a pair are
//                             order-dependent (vMatchedDistance / vnMatches21 stealing, :488, :513-517), the pairs are not.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "window.cuh"

namespace orbm {

__global__ void __launch_bounds__(1024) grid_build_batch_kernel(const orbx_keypoint* __restrict__ keys, const int32_t* __restrict__ counts,
                                                                int cap, GridParams g, int onlyLevel, int32_t* __restrict__ cellOf,
                                                                int32_t* __restrict__ cellStart, int32_t* __restrict__ items) {
    __shared__ int cnt[GRID_CELLS + 1];
    __shared__ int fill[GRID_CELLS];
    __shared__ int warpSum[32];
    const int f = blockIdx.x;
    const orbx_keypoint* keysUn = keys + (size_t)f * cap;
    const int N = min(counts[f], cap);
    cellOf += (size_t)f * cap; items += (size_t)f * cap; cellStart += (size_t)f * (GRID_CELLS + 1);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int c = tid; c <= GRID_CELLS; c += blockDim.x) cnt[c] = 0;
    __syncthreads();
    for (int i = tid; i < N; i += blockDim.x) {
        const int px = (int)roundf(__fmul_rn(__fsub_rn(keysUn[i].x, g.minX), g.invW));       // :507-509
        const int py = (int)roundf(__fmul_rn(__fsub_rn(keysUn[i].y, g.minY), g.invH));
        int c = -1;
        // onlyLevel >= 0: a grid of that pyramid level's key points alone - what GetFeaturesInArea(x, y, r, l, l) returns from the
        // full grid, in the same order (the level gate of :449-479 only drops entries)
        if (!(px < 0 || px >= ORBM_GRID_COLS || py < 0 || py >= ORBM_GRID_ROWS) && (onlyLevel < 0 || keysUn[i].octave == onlyLevel)) {
            c = px * ORBM_GRID_ROWS + py; atomicAdd(&cnt[c], 1);
        }
        cellOf[i] = c;
    }
    __syncthreads();
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

struct InitBatchArgs {
    const orbx_keypoint* keys; const uint32_t* desc; const int32_t* counts; int cap;
    const int32_t* pairs; int npairs;                    // (f1, f2) per pair, or NULL: (p, p + 1)
    const int32_t* cellStart; const int32_t* items;      // per frame, from grid_build_batch_kernel
    GridParams g;
    float windowSize, nnratio; int checkOri;
    float* prev;                                         // [npairs][cap][2] vbPrevMatched in/out, or NULL: F1's own key points
    int32_t* matchedDistance; int32_t* matches21;        // [npairs][cap] scratch
    uint8_t* bins;                                       // [npairs][cap] scratch: rotation bin of an accepted query
    int32_t* matches12; int32_t* nmatches;               // out: [npairs][cap], [npairs]
};

__global__ void __launch_bounds__(128) init_match_batch_kernel(InitBatchArgs a) {
    const int lane = threadIdx.x & 31;
    const int p = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (p >= a.npairs) return;
    const int f1 = a.pairs ? a.pairs[2 * p] : p, f2 = a.pairs ? a.pairs[2 * p + 1] : p + 1;
    const int N1 = min(a.counts[f1], a.cap), N2 = min(a.counts[f2], a.cap);
    const orbx_keypoint* K1 = a.keys + (size_t)f1 * a.cap;
    const orbx_keypoint* K2 = a.keys + (size_t)f2 * a.cap;
    const uint32_t* D1 = a.desc + (size_t)f1 * a.cap * 8;
    const uint32_t* D2 = a.desc + (size_t)f2 * a.cap * 8;
    const int32_t* cellStart = a.cellStart + (size_t)f2 * (GRID_CELLS + 1);
    const int32_t* items = a.items + (size_t)f2 * a.cap;
    int32_t* md = a.matchedDistance + (size_t)p * a.cap;
    int32_t* m21 = a.matches21 + (size_t)p * a.cap;
    int32_t* m12 = a.matches12 + (size_t)p * a.cap;
    uint8_t* bins = a.bins + (size_t)p * a.cap;
    float* prev = a.prev ? a.prev + (size_t)p * a.cap * 2 : nullptr;
    for (int i = lane; i < a.cap; i += 32) { m12[i] = -1; bins[i] = 255; }
    for (int i = lane; i < N2; i += 32) { md[i] = 0x7fffffff; m21[i] = -1; }        // vMatchedDistance(INT_MAX), vnMatches21(-1), :436-437
    __syncwarp();
    int hist = 0;                                        // lane b < 30 holds rotHist[b].size()
    const float factor = __fdiv_rn(1.0f, (float)ORBM_HISTO_LENGTH);
    constexpr uint32_t NONE = 0xffffffffu;               // NONE >> 20 = 0xfff: "no distance", above every real one (<= 256)
    for (int i1 = 0; i1 < N1; i1++) {
        const orbx_keypoint kp1 = K1[i1];
        if (kp1.octave > 0) continue;                    // :451
        WQuery q;
        q.x = prev ? prev[2 * i1] : kp1.x; q.y = prev ? prev[2 * i1 + 1] : kp1.y; q.r = a.windowSize;
        q.minLevel = 0; q.maxLevel = 0;
        const CellRange cr = window_cells(q, a.g);
        if (cr.empty) continue;
        const uint4 d0 = *(const uint4*)(D1 + (size_t)i1 * 8), d1 = *(const uint4*)(D1 + (size_t)i1 * 8 + 4);
        uint32_t best = NONE;                            // (dist << 20) | position in GetFeaturesInArea order
        int bestIdx = -1;
        uint32_t second = 0xfff;
        int pos = 0;
        // The window's cells (ix, y0..y1) are one contiguous item range per grid column; up to 32 columns are fetched at once
        // (lane = column) and their ranges walked as ONE flattened sequence, which is the reference's enumeration order
        // (ix outer, iy inner, insertion order) in ceil(total / 32) steps instead of one dependent step per column.
        for (int c0 = cr.x0; c0 <= cr.x1; c0 += 32) {
            const int ix = c0 + lane;
            int s = 0, len = 0;
            if (ix <= cr.x1) { s = cellStart[ix * ORBM_GRID_ROWS + cr.y0]; len = cellStart[ix * ORBM_GRID_ROWS + cr.y1 + 1] - s; }
            int incl = len;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
            const int total = __shfl_sync(0xffffffffu, incl, 31);
            for (int t0 = 0; t0 < total; t0 += 32) {
                const int t = t0 + lane;
                int col = 0;                             // first column whose inclusive end exceeds t
#pragma unroll
                for (int step = 16; step > 0; step >>= 1) {
                    const int v = __shfl_sync(0xffffffffu, incl, col + step - 1);
                    if (t >= v) col += step;
                }
                col = min(col, 31);
                const int cs = __shfl_sync(0xffffffffu, s, col), cincl = __shfl_sync(0xffffffffu, incl, col), clen = __shfl_sync(0xffffffffu, len, col);
                bool ok = false;
                int idx = -1;
                if (t < total) {
                    idx = items[cs + (t - (cincl - clen))];
                    const orbx_keypoint kp = K2[idx];
                    ok = kp.octave == 0;                 // minLevel = maxLevel = 0, :449-479
                    const float dx = __fsub_rn(kp.x, q.x), dy = __fsub_rn(kp.y, q.y);
                    if (!(fabsf(dx) < q.r && fabsf(dy) < q.r)) ok = false;
                }
                const unsigned m = __ballot_sync(0xffffffffu, ok);
                if (ok) {
                    const uint4 e0 = *(const uint4*)(D2 + (size_t)idx * 8), e1 = *(const uint4*)(D2 + (size_t)idx * 8 + 4);
                    const uint32_t dist = __popc(d0.x ^ e0.x) + __popc(d0.y ^ e0.y) + __popc(d0.z ^ e0.z) + __popc(d0.w ^ e0.w) +
                                          __popc(d1.x ^ e1.x) + __popc(d1.y ^ e1.y) + __popc(d1.z ^ e1.z) + __popc(d1.w ^ e1.w);
                    if (!(__ldcg(&md[idx]) <= (int)dist)) {          // :488
                        const uint32_t key = (dist << 20) | (uint32_t)(pos + __popc(m & ((1u << lane) - 1)));
                        if (key < best) { second = best >> 20; best = key; bestIdx = idx; }
                        else if (dist < second) second = dist;
                    }
                }
                pos += __popc(m);
            }
        }
        if (pos == 0) continue;                          // vIndices2.empty(), :458
        // warp top-2: first strict minimum (the position breaks ties), second smallest distance counted with multiplicity
        uint32_t wb = best;
#pragma unroll
        for (int sft = 16; sft > 0; sft >>= 1) wb = min(wb, __shfl_xor_sync(0xffffffffu, wb, sft));
        if (wb == NONE) continue;                        // every candidate skipped: bestDist stays INT_MAX
        const unsigned owner = __ballot_sync(0xffffffffu, best == wb);
        const int bestIdx2 = __shfl_sync(0xffffffffu, bestIdx, __ffs(owner) - 1);
        uint32_t sec = best == wb ? second : (best >> 20);
#pragma unroll
        for (int sft = 16; sft > 0; sft >>= 1) sec = min(sec, __shfl_xor_sync(0xffffffffu, sec, sft));
        const int bestDist = (int)(wb >> 20);
        const int bestDist2 = sec >= 0xfff ? 0x7fffffff : (int)sec;
        if (bestDist <= ORBM_TH_LOW && (float)bestDist < __fmul_rn((float)bestDist2, a.nnratio)) {      // :505-509
            int bin = -1;
            if (a.checkOri) {                            // :528-540
                float rot = __fsub_rn(kp1.angle, K2[bestIdx2].angle);
                if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
                bin = (int)roundf(__fmul_rn(rot, factor));
                if (bin == ORBM_HISTO_LENGTH) bin = 0;
                if (lane == bin) hist++;
            }
            if (lane == 0) {
                const int prevOwner = __ldcg(&m21[bestIdx2]);
                if (prevOwner >= 0) __stcg(&m12[prevOwner], -1);       // :513-517
                __stcg(&m12[i1], bestIdx2); __stcg(&m21[bestIdx2], i1); __stcg(&md[bestIdx2], bestDist);
                if (bin >= 0) bins[i1] = (uint8_t)bin;
            }
            __syncwarp();
        }
    }
    __syncwarp();
    if (a.checkOri) {                                    // ComputeThreeMaxima (:1679-1723) on the bin sizes, then :548-578
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) {
            const int sz = __shfl_sync(0xffffffffu, hist, i);
            if (sz > max1) { max3 = max2; max2 = max1; max1 = sz; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (sz > max2) { max3 = max2; max2 = sz; ind3 = ind2; ind2 = i; }
            else if (sz > max3) { max3 = sz; ind3 = i; }
        }
        if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) ind3 = -1;
        for (int i = lane; i < N1; i += 32) {
            const int b = bins[i];
            if (b != 255 && b != ind1 && b != ind2 && b != ind3 && __ldcg(&m12[i]) >= 0) __stcg(&m12[i], -1);
        }
        __syncwarp();
    }
    int cnt = 0;
    for (int i0 = 0; i0 < N1; i0 += 32) {
        const int i = i0 + lane;
        const int m = i < N1 ? __ldcg(&m12[i]) : -1;
        if (m >= 0 && prev) { prev[2 * i] = K2[m].x; prev[2 * i + 1] = K2[m].y; }    // :582-584
        cnt += __popc(__ballot_sync(0xffffffffu, m >= 0));
    }
    if (lane == 0) a.nmatches[p] = cnt;
}

// ---- team variant: one block per pair ---------------------------------------------------------------------------------
// The loop of one pair is order-dependent only through vMatchedDistance / vnMatches21 (:488, :513-517); the expensive part of a
// query - walking the grid window and measuring every candidate's descriptor distance, a chain of four dependent global
// loads - is not.  INIT_TEAM enumerator warps therefore prepare the candidate lists of the next INIT_TEAM queries
// ((distance << 20) | index, in GetFeaturesInArea order, in shared memory) while one resolver warp replays the previous
// group's queries in order against the state, which lives in shared memory as well.  Results are identical to the
// one-warp kernel; a pair takes ~1/10 of the time (the one-warp kernel is bound by its serial chain: 0.74 ms per launch no
// matter how few pairs a GPU holds).
constexpr int INIT_TEAM = 8;                   // enumerator warps
constexpr int INIT_LIST = 256;                 // candidates kept per query; longer lists are re-enumerated by the resolver
constexpr int INIT_TEAM_MAX_CAP = 4096;        // key points per frame the shared-memory state is sized for

__host__ __device__ inline size_t init_team_smem(int cap) {
    return (size_t)2 * INIT_TEAM * INIT_LIST * 4 + (size_t)(GRID_CELLS + 4) * 4 + (size_t)cap * (2 + 4 + 4) + 64;
}

struct InitQuery { float x, y, angle; int i1; int count; };     // count: -1 = no query in this slot, > INIT_LIST = overflow

// walks the window of query q over frame 2's level-0 grid in the reference's enumeration order.  STORE: appends
// (dist << 20 | idx) to list; else applies the :488 skip against md and keeps the lane's top-2.  Returns the candidate count.
template <bool STORE>
__device__ __forceinline__ int init_enumerate(const WQuery& q, const CellRange& cr, const uint4 d0, const uint4 d1, const int32_t* __restrict__ cellStart,
                                              const int32_t* __restrict__ items, const orbx_keypoint* __restrict__ K2, const uint32_t* __restrict__ D2,
                                              int lane, uint32_t* __restrict__ list, const unsigned short* __restrict__ md,
                                              uint32_t& best, int& bestIdx, uint32_t& second) {
    int pos = 0;
    for (int c0 = cr.x0; c0 <= cr.x1; c0 += 32) {
        const int ix = c0 + lane;
        int s = 0, len = 0;
        if (ix <= cr.x1) { s = cellStart[ix * ORBM_GRID_ROWS + cr.y0]; len = cellStart[ix * ORBM_GRID_ROWS + cr.y1 + 1] - s; }
        int incl = len;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        for (int t0 = 0; t0 < total; t0 += 32) {
            const int t = t0 + lane;
            int col = 0;                                 // first column whose inclusive end exceeds t
#pragma unroll
            for (int step = 16; step > 0; step >>= 1) {
                const int v = __shfl_sync(0xffffffffu, incl, col + step - 1);
                if (t >= v) col += step;
            }
            col = min(col, 31);
            const int cs = __shfl_sync(0xffffffffu, s, col), cincl = __shfl_sync(0xffffffffu, incl, col), clen = __shfl_sync(0xffffffffu, len, col);
            bool ok = false;
            int idx = -1;
            uint4 e0 = make_uint4(0, 0, 0, 0), e1 = e0;
            if (t < total) {
                idx = items[cs + (t - (cincl - clen))];
                const orbx_keypoint kp = K2[idx];
                e0 = *(const uint4*)(D2 + (size_t)idx * 8); e1 = *(const uint4*)(D2 + (size_t)idx * 8 + 4);   // with the key point, not after it: one round trip less
                ok = kp.octave == 0;                     // minLevel = maxLevel = 0, :449-479
                const float dx = __fsub_rn(kp.x, q.x), dy = __fsub_rn(kp.y, q.y);
                if (!(fabsf(dx) < q.r && fabsf(dy) < q.r)) ok = false;
            }
            const unsigned m = __ballot_sync(0xffffffffu, ok);
            if (ok) {
                const uint32_t dist = __popc(d0.x ^ e0.x) + __popc(d0.y ^ e0.y) + __popc(d0.z ^ e0.z) + __popc(d0.w ^ e0.w) +
                                      __popc(d1.x ^ e1.x) + __popc(d1.y ^ e1.y) + __popc(d1.z ^ e1.z) + __popc(d1.w ^ e1.w);
                const int mypos = pos + __popc(m & ((1u << lane) - 1));
                if (STORE) {
                    if (mypos < INIT_LIST) list[mypos] = (dist << 20) | (uint32_t)idx;
                } else if (!(md[idx] <= dist)) {         // :488
                    const uint32_t key = (dist << 20) | (uint32_t)mypos;
                    if (key < best) { second = best >> 20; best = key; bestIdx = idx; }
                    else if (dist < second) second = dist;
                }
            }
            pos += __popc(m);
        }
    }
    return pos;
}

__global__ void __launch_bounds__((INIT_TEAM + 1) * 32) init_match_team_kernel(InitBatchArgs a) {
    extern __shared__ __align__(16) uint8_t ism[];
    __shared__ InitQuery s_q[2][INIT_TEAM];
    __shared__ int s_n0;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nthr = blockDim.x;
    const int p = blockIdx.x;
    const int f1 = a.pairs ? a.pairs[2 * p] : p, f2 = a.pairs ? a.pairs[2 * p + 1] : p + 1;
    const int N1 = min(a.counts[f1], a.cap), N2 = min(a.counts[f2], a.cap);
    const orbx_keypoint* K1 = a.keys + (size_t)f1 * a.cap;
    const orbx_keypoint* K2 = a.keys + (size_t)f2 * a.cap;
    const uint32_t* D1 = a.desc + (size_t)f1 * a.cap * 8;
    const uint32_t* D2 = a.desc + (size_t)f2 * a.cap * 8;
    const int32_t* cellStart = a.cellStart + (size_t)f2 * (GRID_CELLS + 1);
    const int32_t* items = a.items + (size_t)f2 * a.cap;
    int32_t* m12 = a.matches12 + (size_t)p * a.cap;
    uint8_t* bins = a.bins + (size_t)p * a.cap;
    float* prev = a.prev ? a.prev + (size_t)p * a.cap * 2 : nullptr;
    uint32_t* lists = (uint32_t*)ism;                                        // [2][INIT_TEAM][INIT_LIST]
    int32_t* cs2 = (int32_t*)(ism + (size_t)2 * INIT_TEAM * INIT_LIST * 4);  // frame 2's grid row starts: every query reads two per column
    int32_t* m21 = cs2 + GRID_CELLS + 4;                                     // vnMatches21
    float* ang2 = (float*)(m21 + a.cap);                                     // frame 2's key point angles
    unsigned short* md = (unsigned short*)(ang2 + a.cap);                    // vMatchedDistance (0xffff = INT_MAX)
    if (tid == 0) s_n0 = 0;
    for (int i = tid; i < a.cap; i += nthr) { m12[i] = -1; bins[i] = 255; }
    for (int i = tid; i < N2; i += nthr) { md[i] = 0xffffu; m21[i] = -1; ang2[i] = K2[i].angle; }   // :436-437
    for (int i = tid; i <= GRID_CELLS; i += nthr) cs2[i] = cellStart[i];
    __syncthreads();
    // one past the last level-0 key point of frame 1: later ones are skipped by :451 anyway
    {
        int last = 0;
        for (int i = tid; i < N1; i += nthr) if (K1[i].octave == 0) last = i + 1;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) last = max(last, __shfl_xor_sync(0xffffffffu, last, d));
        if (lane == 0 && last) atomicMax(&s_n0, last);
    }
    __syncthreads();
    const int n0 = s_n0;
    const int ngroups = (n0 + INIT_TEAM - 1) / INIT_TEAM;
    int hist = 0;                                        // resolver: lane b < 30 holds rotHist[b].size()
    const float factor = __fdiv_rn(1.0f, (float)ORBM_HISTO_LENGTH);
    constexpr uint32_t NONE = 0xffffffffu;
    for (int g = 0; g <= ngroups; g++) {
        if (wid > 0) {
            // ---- enumerators: the lists of group g
            const int slot = wid - 1, i1 = g * INIT_TEAM + slot;
            if (g < ngroups) {
                InitQuery iq; iq.i1 = i1; iq.count = -1; iq.x = iq.y = iq.angle = 0.f;
                if (i1 < n0) {
                    const orbx_keypoint kp1 = K1[i1];
                    if (kp1.octave == 0) {               // :451
                        WQuery q;
                        q.x = prev ? prev[2 * i1] : kp1.x; q.y = prev ? prev[2 * i1 + 1] : kp1.y; q.r = a.windowSize;
                        q.minLevel = 0; q.maxLevel = 0;
                        const CellRange cr = window_cells(q, a.g);
                        iq.x = q.x; iq.y = q.y; iq.angle = kp1.angle; iq.count = 0;
                        if (!cr.empty) {
                            const uint4 d0 = *(const uint4*)(D1 + (size_t)i1 * 8), d1 = *(const uint4*)(D1 + (size_t)i1 * 8 + 4);
                            uint32_t b_ = 0, s_ = 0; int bi_ = 0;
                            iq.count = init_enumerate<true>(q, cr, d0, d1, cs2, items, K2, D2, lane,
                                                            lists + ((size_t)(g & 1) * INIT_TEAM + slot) * INIT_LIST, nullptr, b_, bi_, s_);
                        }
                    }
                }
                if (lane == 0) s_q[g & 1][slot] = iq;
            }
        } else if (g > 0) {
            // ---- resolver: the queries of group g - 1, in order, against the state
            const int buf = (g - 1) & 1;
            for (int slot = 0; slot < INIT_TEAM; slot++) {
                const InitQuery iq = s_q[buf][slot];
                if (iq.count <= 0) continue;             // no query, empty window, or vIndices2.empty() (:458)
                uint32_t best = NONE, second = 0xfff;
                int bestIdx = -1;
                if (iq.count <= INIT_LIST) {
                    const uint32_t* list = lists + ((size_t)buf * INIT_TEAM + slot) * INIT_LIST;
                    for (int t = lane; t < iq.count; t += 32) {
                        const uint32_t e = list[t];
                        const uint32_t dist = e >> 20; const int idx = (int)(e & 0xfffffu);
                        if (!(md[idx] <= dist)) {        // :488
                            const uint32_t key = (dist << 20) | (uint32_t)t;
                            if (key < best) { second = best >> 20; best = key; bestIdx = idx; }
                            else if (dist < second) second = dist;
                        }
                    }
                } else {                                 // a window with more candidates than a list holds: walk it here
                    WQuery q; q.x = iq.x; q.y = iq.y; q.r = a.windowSize; q.minLevel = 0; q.maxLevel = 0;
                    const CellRange cr = window_cells(q, a.g);
                    const uint4 d0 = *(const uint4*)(D1 + (size_t)iq.i1 * 8), d1 = *(const uint4*)(D1 + (size_t)iq.i1 * 8 + 4);
                    init_enumerate<false>(q, cr, d0, d1, cs2, items, K2, D2, lane, nullptr, md, best, bestIdx, second);
                }
                // warp top-2: first strict minimum (the position breaks ties), second smallest distance counted with multiplicity
                const uint32_t wb = __reduce_min_sync(0xffffffffu, best);        // (redux.sync: one instruction instead of five shuffle rounds)
                if (wb == NONE) continue;                // every candidate skipped: bestDist stays INT_MAX
                const unsigned owner = __ballot_sync(0xffffffffu, best == wb);
                const int bestIdx2 = __shfl_sync(0xffffffffu, bestIdx, __ffs(owner) - 1);
                const uint32_t sec = __reduce_min_sync(0xffffffffu, best == wb ? second : (best >> 20));
                const int bestDist = (int)(wb >> 20);
                const int bestDist2 = sec >= 0xfff ? 0x7fffffff : (int)sec;
                if (bestDist <= ORBM_TH_LOW && (float)bestDist < __fmul_rn((float)bestDist2, a.nnratio)) {      // :505-509
                    int bin = -1;
                    if (a.checkOri) {                    // :528-540
                        float rot = __fsub_rn(iq.angle, ang2[bestIdx2]);
                        if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
                        bin = (int)roundf(__fmul_rn(rot, factor));
                        if (bin == ORBM_HISTO_LENGTH) bin = 0;
                        if (lane == bin) hist++;
                    }
                    if (lane == 0) {
                        const int prevOwner = m21[bestIdx2];
                        if (prevOwner >= 0) m12[prevOwner] = -1;           // :513-517
                        m12[iq.i1] = bestIdx2; m21[bestIdx2] = iq.i1; md[bestIdx2] = (unsigned short)bestDist;
                        if (bin >= 0) bins[iq.i1] = (uint8_t)bin;
                    }
                    __syncwarp();
                }
            }
        }
        __syncthreads();
    }
    // ---- ComputeThreeMaxima (:1679-1723) on the bin sizes (resolver's registers -> everybody), then :548-578
    __shared__ int s_ind[3];
    if (wid == 0 && a.checkOri) {
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) {
            const int sz = __shfl_sync(0xffffffffu, hist, i);
            if (sz > max1) { max3 = max2; max2 = max1; max1 = sz; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (sz > max2) { max3 = max2; max2 = sz; ind3 = ind2; ind2 = i; }
            else if (sz > max3) { max3 = sz; ind3 = i; }
        }
        if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) ind3 = -1;
        if (lane == 0) { s_ind[0] = ind1; s_ind[1] = ind2; s_ind[2] = ind3; }
    }
    __syncthreads();                                     // (also orders the resolver's global m12 / bins writes before the reads below)
    __shared__ int s_cnt;
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    int cnt = 0;
    for (int i = tid; i < N1; i += nthr) {
        int m = m12[i];
        if (a.checkOri && m >= 0) {
            const int b = bins[i];
            if (b != 255 && b != s_ind[0] && b != s_ind[1] && b != s_ind[2]) { m = -1; m12[i] = -1; }
        }
        if (m >= 0) {
            if (prev) { prev[2 * i] = K2[m].x; prev[2 * i + 1] = K2[m].y; }    // :582-584
            cnt++;
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
    if (lane == 0 && cnt) atomicAdd(&s_cnt, cnt);
    __syncthreads();
    if (tid == 0) a.nmatches[p] = s_cnt;
}

}  // namespace orbm
