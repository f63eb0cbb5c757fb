// Frame::ComputeStereoMatches (reference src/Frame.cc:630-872) on the device, reading the two extractors'
// resident pyramids (the only host-side consumer of mvImagePyramid pixels in the reference, :642,:770,:790,:797).
//   stereo_match_kernel   one warp per left keypoint: row-band candidates (:647-673 - membership is evaluated per
//                         right keypoint instead of materialising vRowIndices; ascending iR keeps the reference's
//                         first-strictly-lowest tie-break), coarse Hamming match (:692-751), 11x11 SAD over +-5 px on
//                         the left keypoint's pyramid level (:756-812), parabola fit and disparity gate (:816-856)
//   stereo_outlier_kernel one block: median of the SAD distances by histogram selection, 1.5*1.4*median cut (:858-871)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "geom.h"
#include "../../include/orbx.h"

namespace orbx {

struct StereoArgs {
    const orbx_keypoint* keysL; const uint32_t* descL; int nL;
    const orbx_keypoint* keysR; const uint32_t* descR; int nR;
    const uint8_t* pyrL; const uint8_t* pyrR;       // the frame's bordered pyramid blocks
    int nlevels;
    unsigned long long pyrOff[ORB_MAX_LEVELS];
    int pitch[ORB_MAX_LEVELS], cols[ORB_MAX_LEVELS];
    float sf[ORB_MAX_LEVELS], isf[ORB_MAX_LEVELS];
    float mbf, maxD;
    float* uRight; float* depth; int* sad;          // [nL]; sad = -1 where unmatched
    // batch of frames (blockIdx.y = frame): every array above is [frames][cap], the counts come from the device
    int cap;                                        // 0: single frame
    const int* nLdev; const int* nRdev;
    unsigned long long pyrFrameStride;
    // vRowIndices (:647-673) of every frame as a CSR over image rows (stereo_rows_kernel), or NULL: scan all right key points
    const int* rowStart; const unsigned short* rowItems; int nRows; int rowItemsCap;
};

constexpr int STEREO_MAX_ROWS = 4096;
constexpr int STEREO_BAND_MAX = 20;                 // rows one right key point can enter with the default settings (8 levels, 1.2): at most 17
// rows one right key point can enter: ceil(y + r) - floor(y - r) + 1 <= 2 * ceil(r) + 3 with r = 2 * scaleFactor[top level];
// configurations above STEREO_BAND_MAX (e.g. scaleFactor 1.5, or 12 levels) scan every right key point instead
__host__ inline int stereo_band_rows(float sfTop) { return 2 * (int)ceilf(2.0f * sfTop) + 3; }

// vRowIndices: right key point iR is a candidate on the image rows floor(y - r) .. ceil(y + r), r = 2 * mvScaleFactors[octave]
// (:658-673).  One block per frame; the order inside a row is free because the match below reduces over (distance, iR).
__global__ void __launch_bounds__(1024) stereo_rows_kernel(const orbx_keypoint* __restrict__ keysR, const int* __restrict__ nRdev, int cap,
                                                           StereoArgs a, int* __restrict__ rowStart, unsigned short* __restrict__ rowItems) {
    __shared__ int cnt[STEREO_MAX_ROWS + 1];
    __shared__ int warpSum[32];
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nR = nRdev ? min(nRdev[f], cap) : cap, nRows = a.nRows;        // single frame: cap = the right key point count
    keysR += (size_t)f * cap; rowStart += (size_t)f * (nRows + 1); rowItems += (size_t)f * a.rowItemsCap;
    for (int r = tid; r <= nRows; r += blockDim.x) cnt[r] = 0;
    __syncthreads();
    for (int iR = tid; iR < nR; iR += blockDim.x) {
        const orbx_keypoint kp = keysR[iR];
        const float r = __fmul_rn(2.0f, a.sf[kp.octave]);
        const int maxr = min((int)ceilf(__fadd_rn(kp.y, r)), nRows - 1), minr = max((int)floorf(__fsub_rn(kp.y, r)), 0);
        for (int y = minr; y <= maxr; y++) atomicAdd(&cnt[y], 1);
    }
    __syncthreads();
    const int per = (nRows + blockDim.x - 1) / blockDim.x;
    const int b = min(tid * per, nRows), e = min(b + per, nRows);
    int s = 0;
    for (int r = b; r < e; r++) s += cnt[r];
    int inc = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
    if (lane == 31) warpSum[wid] = inc;
    __syncthreads();
    int base = 0;
    for (int k = 0; k < wid; k++) base += warpSum[k];
    int run = base + inc - s;
    for (int r = b; r < e; r++) { const int v = cnt[r]; cnt[r] = run; rowStart[r] = run; run += v; }
    if (tid == blockDim.x - 1) rowStart[nRows] = run;
    __syncthreads();
    for (int iR = tid; iR < nR; iR += blockDim.x) {
        const orbx_keypoint kp = keysR[iR];
        const float r = __fmul_rn(2.0f, a.sf[kp.octave]);
        const int maxr = min((int)ceilf(__fadd_rn(kp.y, r)), nRows - 1), minr = max((int)floorf(__fsub_rn(kp.y, r)), 0);
        for (int y = minr; y <= maxr; y++) {
            const int pos = atomicAdd(&cnt[y], 1);
            if (pos < a.rowItemsCap) rowItems[pos] = (unsigned short)iR;       // (the host sizes the list from the real band bound)
        }
    }
}

__global__ void __launch_bounds__(256) stereo_match_kernel(StereoArgs a) {
    const int lane = threadIdx.x & 31;
    const int iL = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    // (the argument block stays in the constant bank: frame offsets go to locals)
    const size_t fr = a.cap ? blockIdx.y : 0, fo = fr * (size_t)a.cap;
    const int nL = a.cap ? min(a.nLdev[fr], a.cap) : a.nL, nR = a.cap ? min(a.nRdev[fr], a.cap) : a.nR;
    const orbx_keypoint* keysL = a.keysL + fo; const uint32_t* descL = a.descL + fo * 8;
    const orbx_keypoint* keysR = a.keysR + fo; const uint32_t* descR = a.descR + fo * 8;
    const uint8_t* pyrL = a.pyrL + fr * a.pyrFrameStride; const uint8_t* pyrR = a.pyrR + fr * a.pyrFrameStride;
    float* uRight = a.uRight + fo; float* depth = a.depth + fo; int* sad = a.sad + fo;
    if (iL >= nL) return;
    const orbx_keypoint kpL = keysL[iL];
    if (lane == 0) { uRight[iL] = -1.0f; depth[iL] = -1.0f; sad[iL] = -1; }     // :634-635
    const int levelL = kpL.octave;
    const float vL = kpL.y, uL = kpL.x;
    const int rowL = (int)vL;                                   // vRowIndices[vL]
    const float minU = __fsub_rn(uL, a.maxD), maxU = uL;       // minD = 0
    if (maxU < 0) return;
    uint32_t dl[8];
#pragma unroll
    for (int k = 0; k < 8; k++) dl[k] = descL[(size_t)iL * 8 + k];
    // ---- coarse match: lowest Hamming distance below TH_HIGH, lowest iR on ties (strict '<' in ascending iR)
    uint32_t best = (100u << 16);                               // (bestDist = TH_HIGH, no index)
    if (a.rowStart) {                                           // the row's candidate list (vRowIndices[vL])
        if (rowL >= 0 && rowL < a.nRows) {
            const int* rs = a.rowStart + fr * (size_t)(a.nRows + 1);
            const unsigned short* ri = a.rowItems + fr * (size_t)a.rowItemsCap;
            for (int k = rs[rowL] + lane, e = rs[rowL + 1]; k < e; k += 32) {
                const int iR = ri[k];
                const orbx_keypoint kpR = keysR[iR];
                if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
                if (!(kpR.x >= minU && kpR.x <= maxU)) continue;
                int d = 0;
#pragma unroll
                for (int w8 = 0; w8 < 8; w8++) d += __popc(dl[w8] ^ descR[(size_t)iR * 8 + w8]);
                if (d < 100) best = min(best, ((uint32_t)d << 16) | (uint32_t)iR);
            }
        }
    } else
    for (int iR = lane; iR < nR; iR += 32) {
        const orbx_keypoint kpR = keysR[iR];
        const float r = __fmul_rn(2.0f, a.sf[kpR.octave]);
        const int maxr = (int)ceilf(__fadd_rn(kpR.y, r)), minr = (int)floorf(__fsub_rn(kpR.y, r));
        if (rowL < minr || rowL > maxr) continue;              // iR is not in vRowIndices[vL]
        if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
        if (!(kpR.x >= minU && kpR.x <= maxU)) continue;
        int d = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) d += __popc(dl[k] ^ descR[(size_t)iR * 8 + k]);
        if (d < 100) best = min(best, ((uint32_t)d << 16) | (uint32_t)iR);
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, s));
    const int bestDist = (int)(best >> 16);
    if (!(bestDist < 75)) return;                               // thOrbDist = (TH_HIGH + TH_LOW) / 2, :639,:757
    const int bestIdxR = (int)(best & 0xffffu);
    // ---- SAD refinement on level `levelL` of both pyramids
    const float uR0 = keysR[bestIdxR].x;
    const float sc = a.isf[levelL];
    const float scaleduL = roundf(__fmul_rn(kpL.x, sc)), scaledvL = roundf(__fmul_rn(kpL.y, sc));
    const float scaleduR0 = roundf(__fmul_rn(uR0, sc));
    const int w = 5, L = 5;
    const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
    if (iniu < 0 || endu >= (float)a.cols[levelL]) return;     // :789
    const int pitch = a.pitch[levelL];
    const uint8_t* imL = pyrL + a.pyrOff[levelL] + (size_t)ORB_EDGE * pitch + ORB_XPAD;
    const uint8_t* imR = pyrR + a.pyrOff[levelL] + (size_t)ORB_EDGE * pitch + ORB_XPAD;
    const int v0 = (int)(scaledvL - w), uL0 = (int)(scaleduL - w), uRc = (int)scaleduR0;
    const int cL = imL[(v0 + w) * pitch + uL0 + w];
    // lane owns window pixels p = lane, lane+32, lane+64, lane+96 (< 121)
    int il[4], pr[4], pc[4];
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const int p = lane + 32 * q;
        pr[q] = p / 11; pc[q] = p - pr[q] * 11;
        il[q] = p < 121 ? (int)imL[(v0 + pr[q]) * pitch + uL0 + pc[q]] - cL : 0;
    }
    int dists[11];
#pragma unroll
    for (int inc = -5; inc <= 5; inc++) {
        const int u0 = uRc + inc - w;
        const int cR = imR[(v0 + w) * pitch + u0 + w];
        int acc = 0;
#pragma unroll
        for (int q = 0; q < 4; q++)
            if (lane + 32 * q < 121) acc += abs(il[q] - ((int)imR[(v0 + pr[q]) * pitch + u0 + pc[q]] - cR));
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
        dists[inc + 5] = acc;
    }
    int bestS = 0x7fffffff, bestinc = 0;
#pragma unroll
    for (int k = 0; k < 11; k++)
        if (dists[k] < bestS) { bestS = dists[k]; bestinc = k - 5; }            // strict '<': first minimum
    if (bestinc == -L || bestinc == L) return;                                  // :816
    float dist1 = 0, dist2 = 0, dist3 = 0;
#pragma unroll
    for (int k = 1; k < 10; k++)
        if (k == bestinc + 5) { dist1 = (float)dists[k - 1]; dist2 = (float)dists[k]; dist3 = (float)dists[k + 1]; }
    const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3),
                                   __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2))));
    if (deltaR < -1 || deltaR > 1) return;
    float bestuR = __fmul_rn(a.sf[levelL], __fadd_rn(__fadd_rn(scaleduR0, (float)bestinc), deltaR));
    float disparity = __fsub_rn(uL, bestuR);
    if (disparity >= 0 && disparity < a.maxD) {
        if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)uL - 0.01); }
        if (lane == 0) { depth[iL] = __fdiv_rn(a.mbf, disparity); uRight[iL] = bestuR; sad[iL] = bestS; }
    }
}

// median-based outlier cut over the matched keypoints of one frame (single block)
// (batch: one block per frame, cap > 0, counts from nLdev)
__global__ void __launch_bounds__(1024) stereo_outlier_kernel(int nL, const int* __restrict__ sad, float* uRight, float* depth,
                                                              int* nMatched, int cap, const int* __restrict__ nLdev) {
    if (cap) {
        const size_t f = blockIdx.x, o = f * (size_t)cap;
        nL = min(nLdev[f], cap); sad += o; uRight += o; depth += o; nMatched += f;
    }
    __shared__ int hist[256];
    __shared__ int sel[4];        // [0] = number of matches, [1] = chosen high byte, [2] = rank inside it, [3] = median
    const int tid = threadIdx.x;
    if (tid < 256) hist[tid] = 0;
    if (tid == 0) sel[0] = 0;
    __syncthreads();
    int cnt = 0;
    for (int i = tid; i < nL; i += blockDim.x) {
        const int s = sad[i];
        if (s >= 0) { atomicAdd(&hist[min(s >> 8, 255)], 1); cnt++; }
    }
    if (cnt) atomicAdd(&sel[0], cnt);
    __syncthreads();
    const int n = sel[0];
    if (n == 0) { if (tid == 0) *nMatched = 0; return; }
    if (tid == 0) {                                   // rank n/2 (0-based) of the ascending order, :861
        int k = n / 2, b = 0;
        while (k >= hist[b]) { k -= hist[b]; b++; }
        sel[1] = b; sel[2] = k;
    }
    __syncthreads();
    const int hb = sel[1];
    __syncthreads();
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < nL; i += blockDim.x) {
        const int s = sad[i];
        if (s >= 0 && min(s >> 8, 255) == hb) atomicAdd(&hist[s & 255], 1);
    }
    __syncthreads();
    if (tid == 0) {
        int k = sel[2], b = 0;
        while (k >= hist[b]) { k -= hist[b]; b++; }
        sel[3] = (hb << 8) | b;
    }
    __syncthreads();
    const float median = (float)sel[3];
    const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), median);
    int kept = 0;
    for (int i = tid; i < nL; i += blockDim.x) {
        const int s = sad[i];
        if (s < 0) continue;
        if ((float)s < thDist) kept++;
        else { uRight[i] = -1.0f; depth[i] = -1.0f; }
    }
    __syncthreads();
    if (tid == 0) sel[0] = 0;
    __syncthreads();
    if (kept) atomicAdd(&sel[0], kept);
    __syncthreads();
    if (tid == 0) *nMatched = sel[0];
}

}  // namespace orbx
