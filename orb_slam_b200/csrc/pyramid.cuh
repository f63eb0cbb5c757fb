// Fused per-level kernel: builds one 128x32 tile of pyramid level l in shared
// memory (INTER_LINEAR fixed-point resize of level l-1, or a copy of the input
// for level 0), then, from that one on-chip tile with its 3-pixel halo,
//   - stores the level (ComputePyramid, reference src/ORBextractor.cc:1342-1377),
//   - stores the FAST-9/16 corner score of every pixel of the detection area
//     (cv::FAST + cornerScore, call sites :976,:985) - 0 where score < minThFAST,
//   - stores the 7x7 sigma-2 fixed-point Gaussian blur
//     (GaussianBlur of the level clone, :1300-1306).
// The level is read from HBM/L2 exactly once by its consumers on this path.
// Integer arithmetic throughout; results are bit-exact with OpenCV 4.13.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "geom.h"

namespace orbx {

constexpr int TW = 128;          // tile width  (pixels)
constexpr int TH = 32;           // tile height
constexpr int HALO = 3;          // FAST ring radius == blur radius
constexpr int TCOL0 = 4;         // shared-memory column of tile pixel x0 (keeps interior word-aligned)
constexpr int TS = 144;          // shared-memory row stride of the level tile
constexpr int TROWS = TH + 2 * HALO;
constexpr int PYR_THREADS = 256;

__device__ __forceinline__ int reflect101(int p, int len) {
    // gfedcb|abcdefgh|gfedcba ; callers guarantee -len < p < 2*len-1
    if (p < 0) p = -p;
    if (p >= len) p = 2 * len - 2 - p;
    return p;
}

__device__ __forceinline__ bool run9(unsigned m16) {
    unsigned x = m16 | (m16 << 16);
    x &= x >> 1;
    x &= x >> 2;
    x &= x >> 4;
    x &= x >> 1;
    return x != 0;
}

// FAST-9/16 score of the pixel at T (shared memory, row stride TS): the largest
// threshold at which it is still a corner, or 0 when that is below minTh.
__device__ __forceinline__ int fast_score(const uint8_t* __restrict__ T, int minTh) {
    const int v = T[0];
    // opposite pairs first: an arc of 9 contains a pixel of every opposite pair
    const int p0 = T[3 * TS], p8 = T[-3 * TS];
    if (abs(p0 - v) <= minTh && abs(p8 - v) <= minTh) return 0;
    const int p4 = T[3], p12 = T[-3];
    if (abs(p4 - v) <= minTh && abs(p12 - v) <= minTh) return 0;
    int e[16];   // v - p_k, clockwise ring as OpenCV numbers it
    e[0] = v - p0;                 e[1] = v - T[1 + 3 * TS];   e[2] = v - T[2 + 2 * TS];  e[3] = v - T[3 + TS];
    e[4] = v - p4;                 e[5] = v - T[3 - TS];       e[6] = v - T[2 - 2 * TS];  e[7] = v - T[1 - 3 * TS];
    e[8] = v - p8;                 e[9] = v - T[-1 - 3 * TS];  e[10] = v - T[-2 - 2 * TS]; e[11] = v - T[-3 - TS];
    e[12] = v - p12;               e[13] = v - T[-3 + TS];     e[14] = v - T[-2 + 2 * TS]; e[15] = v - T[-1 + 3 * TS];
    unsigned mdark = 0, mbright = 0;   // ring pixel darker / brighter than the centre by more than minTh
#pragma unroll
    for (int k = 0; k < 16; k++) {
        mdark |= (unsigned)(e[k] > minTh) << k;
        mbright |= (unsigned)(e[k] < -minTh) << k;
    }
    if (!run9(mdark) && !run9(mbright)) return 0;
    // score+1 = max over the 16 arcs of 9 of max( min(v-p), min(p-v) )  (cornerScore<16>).
    // Both halves are written as MIN networks over e and -e.  (The equivalent max-network form,
    // best = max(min9(e), -max9(e)), is miscompiled by nvcc 12.9 -O3 for sm_100a: verified on a
    // B200 against the same source built with -Xcicc -O0; see DESIGN.md "toolchain notes".)
    int f[16];
#pragma unroll
    for (int k = 0; k < 16; k++) f[k] = -e[k];
    int a2[16], b2[16], a4[16], b4[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { a2[k] = min(e[k], e[(k + 1) & 15]); b2[k] = min(f[k], f[(k + 1) & 15]); }
#pragma unroll
    for (int k = 0; k < 16; k++) { a4[k] = min(a2[k], a2[(k + 2) & 15]); b4[k] = min(b2[k], b2[(k + 2) & 15]); }
    int best = -256;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const int a9 = min(min(a4[k], a4[(k + 4) & 15]), e[(k + 8) & 15]);
        const int b9 = min(min(b4[k], b4[(k + 4) & 15]), f[(k + 8) & 15]);
        best = max(best, max(a9, b9));
    }
    const int score = best - 1;
    return score >= minTh ? score : 0;
}

struct PyrArgs {
    const uint8_t* src;          // level l-1 planes (frame 0), or the input images for level 0
    unsigned long long srcFrameStride;
    int srcPitch;                // bytes per source row
    int srcX0;                   // byte offset of source column 0 in a row
    int sw, sh;
    uint8_t* pyr;                // level l bordered plane of frame 0 (points at plane start)
    unsigned long long pyrFrameStride;
    uint8_t* blur;               // level l blur plane of frame 0
    uint8_t* score;              // level l score plane of frame 0
    unsigned long long planeFrameStride;
    int pitch;                   // pitch of level l planes
    int w, h;
    const ResizeTap* xtab;       // w entries (null for level 0)
    const ResizeTap* ytab;       // h entries
    int srcTileCols, srcTileRows;   // staging capacity (level > 0)
    int minTh;
    int* err;                    // device error flag
};

template <bool LEVEL0>
__global__ void __launch_bounds__(PYR_THREADS) pyr_level_kernel(PyrArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t* T = smem;                                   // TROWS x TS level tile with halo
    unsigned short* Hs = (unsigned short*)(smem + TROWS * TS);          // TROWS x TW horizontal blur sums
    uint8_t* S = smem + TROWS * TS + TROWS * TW * 2;     // source staging, later reused for output staging

    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
    const int frame = blockIdx.z;
    const int w = a.w, h = a.h;
    const uint8_t* src = a.src + (size_t)frame * a.srcFrameStride;

    if (LEVEL0) {
        // level 0 is the caller's image: copy with REFLECT_101 halo (copyMakeBorder, :1372)
        for (int i = tid; i < TROWS * (TW + 2 * HALO); i += PYR_THREADS) {
            const int ty = i / (TW + 2 * HALO), tx = i - ty * (TW + 2 * HALO);
            const int x = x0 - HALO + tx, y = y0 - HALO + ty;
            uint8_t v = 0;
            if (x < w + HALO && y < h + HALO)
                v = __ldg(src + (size_t)reflect101(y, h) * a.srcPitch + a.srcX0 + reflect101(x, w));
            T[ty * TS + TCOL0 - HALO + tx] = v;
        }
    } else {
        // stage the source rows/columns this tile touches (16-byte vector loads), then interpolate
        const int ylo = max(0, y0 - HALO), yhi = min(h - 1, y0 + TH + HALO - 1);
        const int xlo = max(0, x0 - HALO), xhi = min(w - 1, x0 + TW + HALO - 1);
        const int sy_lo = a.ytab[ylo].s0, sy_hi = a.ytab[yhi].s1;
        const int sx_lo = a.xtab[xlo].s0 & ~15, sx_hi = a.xtab[xhi].s1;
        const int scols = a.srcTileCols;                       // multiple of 16
        const int nvec = (sx_hi - sx_lo) / 16 + 1;
        const int nrows = sy_hi - sy_lo + 1;
        if (nrows > a.srcTileRows || nvec * 16 > scols) { if (tid == 0) atomicExch(a.err, 201); return; }   // block-uniform
        for (int i = tid; i < nrows * nvec; i += PYR_THREADS) {
            const int r = i / nvec, c = i - r * nvec;
            const uint4 v = __ldg((const uint4*)(src + (size_t)(sy_lo + r) * a.srcPitch + a.srcX0 + sx_lo + 16 * c));
            *(uint4*)(S + r * scols + 16 * c) = v;
        }
        __syncthreads();
        for (int i = tid; i < TROWS * (TW + 2 * HALO); i += PYR_THREADS) {
            const int ty = i / (TW + 2 * HALO), tx = i - ty * (TW + 2 * HALO);
            const int x = x0 - HALO + tx, y = y0 - HALO + ty;
            uint8_t v = 0;
            if (x < w + HALO && y < h + HALO) {
                const ResizeTap tx_ = a.xtab[reflect101(x, w)];
                const ResizeTap ty_ = a.ytab[reflect101(y, h)];
                const uint8_t* r0 = S + (ty_.s0 - sy_lo) * scols - sx_lo;
                const uint8_t* r1 = S + (ty_.s1 - sy_lo) * scols - sx_lo;
                const int h0 = r0[tx_.s0] * tx_.a0 + r0[tx_.s1] * tx_.a1;      // HResizeLinear, int32
                const int h1 = r1[tx_.s0] * tx_.a0 + r1[tx_.s1] * tx_.a1;
                v = (uint8_t)((((ty_.a0 * (h0 >> 4)) >> 16) + ((ty_.a1 * (h1 >> 4)) >> 16) + 2) >> 2);   // VResizeLinear 8u
            }
            T[ty * TS + TCOL0 - HALO + tx] = v;
        }
    }
    __syncthreads();

    const int vx = (tid & 7) * 16, vy = tid >> 3;            // this thread's 16-pixel output segment
    const bool vstore = (y0 + vy < h) && (x0 + vx < w);
    const size_t rowOff = (size_t)(y0 + vy) * a.pitch + ORB_XPAD + x0 + vx;

    // ---- level plane -------------------------------------------------------
    if (vstore) {
        const uint32_t* t = (const uint32_t*)(T + (vy + HALO) * TS + TCOL0 + vx);
        uint4 v = make_uint4(t[0], t[1], t[2], t[3]);
        *(uint4*)(a.pyr + (size_t)frame * a.pyrFrameStride + (size_t)ORB_EDGE * a.pitch + rowOff) = v;
    }

    // ---- horizontal blur sums (exact, <= 16 bit) ----------------------------
    for (int i = tid; i < TROWS * TW; i += PYR_THREADS) {
        const int r = i / TW, c = i - r * TW;
        const uint8_t* p = T + r * TS + TCOL0 - HALO + c;
        Hs[i] = (unsigned short)(18 * (p[0] + p[6]) + 34 * (p[1] + p[5]) + 48 * (p[2] + p[4]) + 56 * p[3]);
    }

    // ---- FAST score (reads T only) ------------------------------------------
    uint8_t* O = S;   // TH x TW output staging; the source staging is dead by now
    __syncthreads();  // all reads of S done (level>0), Hs complete
    for (int i = tid; i < TH * TW; i += PYR_THREADS) {
        const int r = i / TW, c = i - r * TW;
        const int x = x0 + c, y = y0 + r;
        int s = 0;
        if (x >= ORB_EDGE && x < w - ORB_EDGE && y >= ORB_EDGE && y < h - ORB_EDGE)
            s = fast_score(T + (r + HALO) * TS + TCOL0 + c, a.minTh);
        O[i] = (uint8_t)s;
    }
    __syncthreads();
    if (vstore) {
        const uint4 v = *(const uint4*)(O + vy * TW + vx);
        *(uint4*)(a.score + (size_t)frame * a.planeFrameStride + rowOff) = v;
    }
    __syncthreads();

    // ---- vertical blur pass ---------------------------------------------------
    for (int i = tid; i < TH * TW; i += PYR_THREADS) {
        const int r = i / TW, c = i - r * TW;
        const unsigned short* q = Hs + r * TW + c;
        const unsigned v = 18u * (q[0] + q[6 * TW]) + 34u * (q[TW] + q[5 * TW]) + 48u * (q[2 * TW] + q[4 * TW]) + 56u * q[3 * TW];
        O[i] = (uint8_t)((v + 32768u) >> 16);
    }
    __syncthreads();
    if (vstore) {
        const uint4 v = *(const uint4*)(O + vy * TW + vx);
        *(uint4*)(a.blur + (size_t)frame * a.planeFrameStride + rowOff) = v;
    }
}

// BORDER_REFLECT_101 frame of 19 pixels around every level (copyMakeBorder,
// :1364-1365,:1372-1373).  One thread per border pixel.
struct BorderArgs {
    uint8_t* pyr;                 // frame 0 pyramid block
    unsigned long long pyrFrameStride;
    int nlevels;
    int itemOff[ORB_MAX_LEVELS + 1];   // prefix sum of border pixels per level
    unsigned long long pyrOff[ORB_MAX_LEVELS];
    int w[ORB_MAX_LEVELS], h[ORB_MAX_LEVELS], pitch[ORB_MAX_LEVELS];
};

__global__ void border_kernel(BorderArgs a) {
    const int item = blockIdx.x * blockDim.x + threadIdx.x;
    if (item >= a.itemOff[a.nlevels]) return;
    int l = 0;
    while (item >= a.itemOff[l + 1]) l++;
    int i = item - a.itemOff[l];
    const int w = a.w[l], h = a.h[l], bw = w + 2 * ORB_EDGE;
    int bx, by;   // coordinates in the bordered plane
    if (i < ORB_EDGE * bw) { by = i / bw; bx = i - by * bw; }
    else if ((i -= ORB_EDGE * bw) < ORB_EDGE * bw) { by = i / bw; bx = i - by * bw; by += ORB_EDGE + h; }
    else {
        i -= ORB_EDGE * bw;
        by = i / (2 * ORB_EDGE); bx = i - by * (2 * ORB_EDGE);
        by += ORB_EDGE;
        if (bx >= ORB_EDGE) bx += w;
    }
    uint8_t* plane = a.pyr + (size_t)blockIdx.y * a.pyrFrameStride + a.pyrOff[l];
    const int sx = reflect101(bx - ORB_EDGE, w), sy = reflect101(by - ORB_EDGE, h);
    const int xbase = ORB_XPAD - ORB_EDGE;
    plane[(size_t)by * a.pitch[l] + xbase + bx] = plane[(size_t)(sy + ORB_EDGE) * a.pitch[l] + ORB_XPAD + sx];
}

}  // namespace orbx
