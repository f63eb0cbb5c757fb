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
//
// Instruction economy (the kernel is bound by the integer ALU pipe, not by HBM):
// every thread works on 4 horizontally adjacent pixels held as two u16x2 registers
// (pixels c,c+2 and c+1,c+3), shared memory is read as 32-bit words and split with
// PRMT, and the 9-of-16 arc order statistics of FAST run on VIMNMX3.U16x2 (DPX).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "geom.h"
#include "pdl.cuh"

namespace orbx {

constexpr int TW = 128;          // tile width  (pixels)
constexpr int TH_BATCH = 32;     // tile height: throughput variant
constexpr int TH_LATENCY = 16;   // tile height of the few-frames variant: twice the CTAs, half the serial work in each
constexpr int PYR_LATENCY_FRAMES = 2;   // calls with at most this many frames take the few-frames variant
constexpr int HALO = 3;          // FAST ring radius == blur radius
constexpr int TCOL0 = 4;         // shared-memory byte of tile pixel x0 in a T row (keeps 4-pixel groups word-aligned)
constexpr int TS = 144;          // shared-memory row stride of the level tile (bytes)
constexpr int TBYTES = 136;      // T bytes written per row: byte b <-> level column x0 - 4 + b
constexpr int HQS = 136;         // u16 stride of the horizontally resized staging rows
constexpr int PYR_THREADS = 256;
constexpr int PYR_THREADS_LATENCY = 512;     // calls of a frame or two (see pyr_level_kernel)

__device__ __forceinline__ int reflect101(int p, int len) {
    // gfedcb|abcdefgh|gfedcba ; callers guarantee -len < p < 2*len-1
    if (p < 0) p = -p;
    if (p >= len) p = 2 * len - 2 - p;
    return p;
}

// ---- packed helpers ---------------------------------------------------------
// A 32-bit word w of four pixels (b0..b3) splits into E = (b0,b2) and O = (b1,b3) as u16x2 (zero-extended bytes);
// cross(lo, hi) of two such registers is (lo.hi16, hi.lo16): the stride-2 pair straddling two words.
__device__ __forceinline__ uint32_t pk_even(uint32_t w) { return __byte_perm(w, 0u, 0x4240); }
__device__ __forceinline__ uint32_t pk_odd(uint32_t w) { return __byte_perm(w, 0u, 0x4341); }
__device__ __forceinline__ uint32_t pk_cross(uint32_t lo, uint32_t hi) { return __byte_perm(lo, hi, 0x1412); }

// the eight stride-2 pixel pairs around a 4-pixel group at column c that FAST and the 7-tap blur need from one row
struct RowPairs {
    uint32_t ol;   // (c-3, c-1)
    uint32_t x3;   // (c-2, c  )
    uint32_t x1;   // (c-1, c+1)
    uint32_t em;   // (c  , c+2)
    uint32_t om;   // (c+1, c+3)
    uint32_t x2;   // (c+2, c+4)
    uint32_t x4;   // (c+3, c+5)
    uint32_t er;   // (c+4, c+6)
};

// row points at the word holding pixels c-4..c-1
__device__ __forceinline__ RowPairs row_pairs(const uint32_t* __restrict__ row) {
    const uint32_t L = row[0], M = row[1], R = row[2];
    const uint32_t el = pk_even(L), ol = pk_odd(L), em = pk_even(M), om = pk_odd(M), er = pk_even(R), orr = pk_odd(R);
    RowPairs p;
    p.ol = ol; p.em = em; p.om = om; p.er = er;
    p.x1 = pk_cross(ol, om); p.x2 = pk_cross(em, er); p.x3 = pk_cross(el, em); p.x4 = pk_cross(om, orr);
    return p;
}

// max over the 16 arcs of 9 contiguous ring pixels of the arc minimum (MAXOFMIN) or min of the arc maximum,
// lane-wise on u16x2, in 47 two-input comparisons: arcs 2i and 2i+1 share the 8 pixels 2i+1..2i+8, and
//   max( min(m8, P[2i]), min(m8, P[2i+9]) ) = min( m8, max(P[2i], P[2i+9]) ),
// so only the eight odd-aligned windows m8[i] = op(P[2i+1..2i+8]) are built (log-doubling, 24 comparisons), each is
// closed with one pre-combined end pixel (16), and the eight results are reduced (7).  Two-input VIMNMX.U16x2 issues
// at twice the rate of VIMNMX3 on sm_100 (tools/pipebench.cu), so comparisons - not instructions - are the cost.
template <bool MAXOFMIN>
__device__ __forceinline__ uint32_t arc9_extreme(const uint32_t (&P)[16]) {
#define ARC_OP(x, y) (MAXOFMIN ? __vminu2((x), (y)) : __vmaxu2((x), (y)))
#define ARC_RO(x, y) (MAXOFMIN ? __vmaxu2((x), (y)) : __vminu2((x), (y)))
    uint32_t m2[8], m4[8], r[8];
#pragma unroll
    for (int i = 0; i < 8; i++) m2[i] = ARC_OP(P[2 * i + 1], P[(2 * i + 2) & 15]);
#pragma unroll
    for (int i = 0; i < 8; i++) m4[i] = ARC_OP(m2[i], m2[(i + 1) & 7]);
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t m8 = ARC_OP(m4[i], m4[(i + 2) & 7]);              // pixels 2i+1 .. 2i+8
        const uint32_t e = ARC_RO(P[2 * i], P[(2 * i + 9) & 15]);          // the two possible ninth pixels
        r[i] = ARC_OP(m8, e);
    }
    const uint32_t a = ARC_RO(ARC_RO(r[0], r[1]), ARC_RO(r[2], r[3]));
    const uint32_t b = ARC_RO(ARC_RO(r[4], r[5]), ARC_RO(r[6], r[7]));
    return ARC_RO(a, b);
#undef ARC_OP
#undef ARC_RO
}

// cornerScore<16> of OpenCV for two pixels at once (u16x2 lanes): the largest threshold at which the centre is still
// a FAST-9/16 corner = max( max_arcs min9(p) - v , v - min_arcs max9(p) ) - 1.  Returns lanes holding score+257.
__device__ __forceinline__ uint32_t fast_score_pair(const uint32_t (&P)[16], uint32_t V) {
    const uint32_t hi = arc9_extreme<true>(P);     // brightest "all 9 brighter" bound
    const uint32_t lo = arc9_extreme<false>(P);    // darkest  "all 9 darker" bound
    const uint32_t bias = 0x01000100u;
    return __vmaxu2(hi + bias - V, V + bias - lo);
}

// FAST score of the 4 pixels of group `g` on tile row r (0..TH-1), from the pair planes of the 7 tile rows r..r+6
// (PA[row][g] = {x1, x2, em, om}, PB[row][g] = {x4, x3, ol, er}: each row's stride-2 pixel pairs are split out of
// the level tile once, by the row pass, instead of once per ring row that needs them).  The field order puts what the ring
// rows dy = +-2 need (x1, x2 / x4, x3) and what the centre row needs of PA (em, om) into one 64-bit half each, so five of the
// twelve reads are 64-bit: 30 shared-memory wavefronts per warp and row instead of 48 - the kernel's L1 data pipe is as busy
// as its ALU pipe (64 % against 70 %, profiles/r2n_ncu_full_b512.md), and two thirds of its shared-memory reads are these.
// Result: four bytes (pixel order) of max(0, score + 1 - minTh) - the score plane's encoding.
__device__ __forceinline__ uint32_t fast_score4(const uint4* __restrict__ a, const uint4* __restrict__ b, int minTh) {   // a, b: PA / PB at [r][g]
    const uint4 a0 = a[0], a6 = a[6 * 32];                       // dy = -3, +3: x1 x2 em om
    const uint2 a1 = *(const uint2*)(a + 32), a5 = *(const uint2*)(a + 5 * 32);      // dy = -2, +2: x1 x2 ...
    const uint2 b1 = *(const uint2*)(b + 32), b5 = *(const uint2*)(b + 5 * 32);      //              ... x4 x3
    const uint4 b2 = b[2 * 32], b4 = b[4 * 32];                  // dy = -1, +1: x4 x3 ol er
    const uint4 b3 = b[3 * 32];                                  // centre row: x4 x3 ol er
    const uint2 a3 = *((const uint2*)(a + 3 * 32) + 1);          //           + em om
    // ring k -> (dx,dy): 0:(0,3) 1:(1,3) 2:(2,2) 3:(3,1) 4:(3,0) 5:(3,-1) 6:(2,-2) 7:(1,-3) 8:(0,-3) 9:(-1,-3)
    //                    10:(-2,-2) 11:(-3,-1) 12:(-3,0) 13:(-3,1) 14:(-2,2) 15:(-1,3)
    // pair A = pixels (c, c+2): offset dx -> -3:ol -2:x3 -1:x1 0:em 1:om 2:x2 3:x4
    // pair B = pixels (c+1, c+3):          -3:x3 -2:x1 -1:em 0:om 1:x2 2:x4 3:er
    // fields: PA = (x: x1, y: x2, z: em, w: om), PB = (x: x4, y: x3, z: ol, w: er)
    const uint32_t PAr[16] = {a6.z, a6.w, a5.y, b4.x, b3.x, b2.x, a1.y, a0.w,
                              a0.z, a0.x, b1.y, b2.z, b3.z, b4.z, b5.y, a6.x};
    const uint32_t PBr[16] = {a6.w, a6.y, b5.x, b4.w, b3.w, b2.w, b1.x, a0.y,
                              a0.w, a0.z, a1.x, b2.y, b3.y, b4.y, a5.x, a6.z};
    const uint32_t tA = fast_score_pair(PAr, a3.x);
    const uint32_t tB = fast_score_pair(PBr, a3.y);
    // lanes hold score + 257; the plane stores max(0, score + 1 - minTh): 0 <=> not a corner at minTh
    const uint32_t D = 0x01000100u + (uint32_t)minTh * 0x00010001u;
    const uint32_t uA = __vmaxu2(tA, D) - D, uB = __vmaxu2(tB, D) - D;
    return __byte_perm(uA, uB, 0x6240);                  // bytes: pixel c, c+1, c+2, c+3
}

// 7-tap blur row sums (kernel 18,34,48,56,48,34,18) of the 4 pixels of a group, from one row's pairs:
// A = (H(c), H(c+2)), B = (H(c+1), H(c+3)) as exact u16x2 (max 255*256 = 65280).
__device__ __forceinline__ void hblur4(const RowPairs& p, uint32_t& A, uint32_t& B) {
    A = 18u * (p.ol + p.x4) + 34u * (p.x3 + p.x2) + 48u * (p.x1 + p.om) + 56u * p.em;
    B = 18u * (p.x3 + p.er) + 34u * (p.x1 + p.x4) + 48u * (p.em + p.x2) + 56u * p.om;
}

struct TapX { unsigned short o0, o1; short a0, a1; };   // source byte offsets inside the staged row + 11-bit weights

constexpr int PYR_MAX_TILES_X = 64, PYR_MAX_TILES_Y = 160;    // 8192 x 5120 px; larger levels read the tap tables instead

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
    int haveTiles;               // tileX / tileY are filled in (the level has few enough tiles)
    uint32_t tileX[PYR_MAX_TILES_X];   // per tile column: first staged source column (16-aligned) | last source column << 16
    uint32_t tileY[PYR_MAX_TILES_Y];   // per tile row: first | last source row
};

// ---- the phases of a tile once the level tile T (TROWS x TS bytes, halo included) is complete in shared memory ----

// level plane: 16 pixels per thread
template <int TH>
__device__ __forceinline__ void tile_store_plane(const PyrArgs& a, const uint8_t* __restrict__ T, int frame, int x0, int y0, int tid) {
    const int vx = (tid & 7) * 16, vy = tid >> 3;
    if (vy < TH && (y0 + vy < a.h) && (x0 + vx < a.w)) {
        const uint32_t* t = (const uint32_t*)(T + (vy + HALO) * TS + TCOL0 + vx);
        const uint4 v = make_uint4(t[0], t[1], t[2], t[3]);
        *(uint4*)(a.pyr + (size_t)frame * a.pyrFrameStride + (size_t)(ORB_EDGE + y0 + vy) * a.pitch + ORB_XPAD + x0 + vx) = v;
    }
}

// row pass: one warp per T row, one 4-pixel group per lane: split the row into its stride-2 pixel pairs
// (kept in the pair planes for FAST) and form the horizontal blur sums
template <int TH, int NT = PYR_THREADS>
__device__ __forceinline__ void tile_row_pass(const uint8_t* __restrict__ T, unsigned short* __restrict__ Hs, uint4* __restrict__ PA,
                                              uint4* __restrict__ PB, int h, int y0, int wid, int lane) {
    constexpr int TROWS = TH + 2 * HALO;
    const int rEnd = min(TROWS, h + 2 * HALO - y0);       // rows below the image + halo feed nothing that is stored
    for (int r = wid; r < rEnd; r += NT / 32) {
        const RowPairs p = row_pairs((const uint32_t*)(T + r * TS) + lane);
        PA[r * 32 + lane] = make_uint4(p.x1, p.x2, p.em, p.om);
        PB[r * 32 + lane] = make_uint4(p.x4, p.x3, p.ol, p.er);
        uint32_t A, B;
        hblur4(p, A, B);
        *(uint2*)(Hs + r * TW + 4 * lane) = make_uint2(A, B);
    }
}

// FAST score: warp `wid` owns tile rows RPW*wid..RPW*wid+RPW-1, lane = 4-pixel group
template <int TH, int NT = PYR_THREADS>
__device__ __forceinline__ void tile_fast_pass(const PyrArgs& a, const uint4* __restrict__ PA, const uint4* __restrict__ PB, int frame,
                                               int x0, int y0, int wid, int lane) {
    constexpr int RPW = TH / (NT / 32);
    const int w = a.w, h = a.h;
    const int gx = x0 + 4 * lane;                        // level column of the group's first pixel
    if (gx >= w) return;                                 // (no barrier inside this pass)
    // everything that does not change from row to row is formed once: the plane pointers and the store pointer advance by a
    // row per step, the column part of the detection-area test and the mask of a group that straddles the area's edge are fixed
    const int r0 = RPW * wid;
    const uint4* pa = PA + r0 * 32 + lane;
    const uint4* pb = PB + r0 * 32 + lane;
    uint8_t* sp = a.score + (size_t)frame * a.planeFrameStride + (size_t)(y0 + r0) * a.pitch + ORB_XPAD + gx;
    const bool colIn = gx + 3 >= ORB_EDGE && gx < w - ORB_EDGE;
    uint32_t colMask = 0;
#pragma unroll
    for (int j = 0; j < 4; j++)
        if (gx + j >= ORB_EDGE && gx + j < w - ORB_EDGE) colMask |= 0xffu << (8 * j);
    const int rows = min(RPW, h - (y0 + r0));            // tile rows of this warp inside the image
    const int yLo = ORB_EDGE - (y0 + r0), yHi = h - ORB_EDGE - (y0 + r0);    // k in [yLo, yHi) is inside the detection area
#pragma unroll 1
    for (int k = 0; k < rows; k++) {
        uint32_t s = 0;
        if (colIn && k >= yLo && k < yHi) s = fast_score4(pa, pb, a.minTh) & colMask;
        *(uint32_t*)sp = s;
        sp += a.pitch; pa += 32; pb += 32;
    }
}

// vertical blur pass: warp `wid` owns tile rows RPW*wid..RPW*wid+RPW-1.  The RPW + 6 row sums a lane needs per pixel are
// paired vertically, P[j] = (H[j], H[j+1]) as u16x2, so that one DP2A applies two taps:
// out[k] = (18,34).P[k] + (48,56).P[k+2] + (48,34).P[k+4] + (0,18).P[k+5]
template <int TH, int NT = PYR_THREADS>
__device__ __forceinline__ void tile_vblur_pass(const PyrArgs& a, const unsigned short* __restrict__ Hs, int frame, int x0, int y0,
                                                int wid, int lane) {
    constexpr int RPW = TH / (NT / 32);
    const int w = a.w, h = a.h;
    const int gx = x0 + 4 * lane;
    const unsigned short* col = Hs + 4 * lane;
    uint2 v[RPW + 6];
#pragma unroll
    for (int j = 0; j < RPW + 6; j++) v[j] = *(const uint2*)(col + (RPW * wid + j) * TW);     // .x = (H(c), H(c+2)), .y = (H(c+1), H(c+3))
    uint32_t P[RPW + 5][4];                                 // pixel order c, c+1, c+2, c+3
#pragma unroll
    for (int j = 0; j < RPW + 5; j++) {
        P[j][0] = __byte_perm(v[j].x, v[j + 1].x, 0x5410); P[j][2] = __byte_perm(v[j].x, v[j + 1].x, 0x7632);
        P[j][1] = __byte_perm(v[j].y, v[j + 1].y, 0x5410); P[j][3] = __byte_perm(v[j].y, v[j + 1].y, 0x7632);
    }
#pragma unroll
    for (int k = 0; k < RPW; k++) {
        const int y = y0 + RPW * wid + k;
        uint32_t o[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            uint32_t acc = __dp2a_lo(P[k][q], 18u | (34u << 8), 32768u);
            acc = __dp2a_lo(P[k + 2][q], 48u | (56u << 8), acc);
            acc = __dp2a_lo(P[k + 4][q], 48u | (34u << 8), acc);
            acc = __dp2a_lo(P[k + 5][q], 18u << 8, acc);
            o[q] = acc >> 16;
        }
        if (y < h && gx < w)
            *(uint32_t*)(a.blur + (size_t)frame * a.planeFrameStride + (size_t)y * a.pitch + ORB_XPAD + gx) =
                o[0] | (o[1] << 8) | (o[2] << 16) | (o[3] << 24);
    }
}

// ---- building the level tile T of a resized level from staged source rows S (dense rows of `scols` bytes) ----

// per-tile tap tables: T byte b <-> level column x0 - 4 + b (b = 0 and b = 135 unused -> zero weights)
template <int TH>
__device__ __forceinline__ void tile_build_taps(const PyrArgs& a, TapX* __restrict__ tapx, TapX* __restrict__ tapy, int x0, int y0,
                                                int sx_lo, int sy_lo, int tid) {
    constexpr int TROWS = TH + 2 * HALO;
    const int w = a.w, h = a.h;
    if (tid < TBYTES) {
        const int x = x0 - 4 + tid;
        TapX t; t.o0 = 0; t.o1 = 0; t.a0 = 0; t.a1 = 0;
        if (tid >= 1 && tid <= TW + 2 * HALO && x < w + HALO) {
            const ResizeTap s = a.xtab[reflect101(x, w)];
            t.o0 = (unsigned short)(s.s0 - sx_lo); t.o1 = (unsigned short)(s.s1 - sx_lo); t.a0 = s.a0; t.a1 = s.a1;
        }
        tapx[tid] = t;
    } else if (tid >= 192 && tid < 192 + TROWS) {
        const int ty = tid - 192, y = y0 - HALO + ty;
        TapX t; t.o0 = 0; t.o1 = 0; t.a0 = 0; t.a1 = 0;
        if (y < h + HALO) {
            const ResizeTap s = a.ytab[reflect101(y, h)];
            t.o0 = (unsigned short)(s.s0 - sy_lo); t.o1 = (unsigned short)(s.s1 - sy_lo); t.a0 = s.a0; t.a1 = s.a1;
        }
        tapy[ty] = t;
    }
}

// Both resize passes: thread = (4-column group q, row slot); 34 groups x 7 slots = 238 active threads, so the
// column taps of the horizontal pass are decoded once per thread and stay in registers.
// horizontal pass (HResizeLinear, int32): Hq[r][b] = (S[r][o0]*a0 + S[r][o0+1]*a1) >> 4.
// o1 == o0 + 1 whenever a1 != 0 (the x taps collapse the weight, not the index, at the borders).
// A group's four tap pairs lie within 8 consecutive source bytes (scale < 2): fetch that window as three
// aligned words + two funnel shifts, pick each (S[o], S[o+1]) pair with a per-thread constant PRMT selector
// and weight it with one DP2A.
template <int NT = PYR_THREADS>
__device__ __forceinline__ void tile_hresize(const PyrArgs& a, const uint8_t* __restrict__ S, int scols, int nrows,
                                             const TapX* __restrict__ tapx, unsigned short* __restrict__ Hq, int x0, int tid) {
    constexpr int SLOTS = NT / (TBYTES / 4);             // row slots: 7 at 256 threads, 15 at 512
    const int q = tid % (TBYTES / 4), slot = tid / (TBYTES / 4);
    const bool live = slot < SLOTS && x0 - 4 + 4 * q < a.w + HALO;     // groups right of the image + halo are never read
    if (!live) return;
    int o[4];
    uint32_t wt[4];
    int base = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const uint2 tp = *(const uint2*)(tapx + 4 * q + j);
        o[j] = (int)(tp.x & 0xffffu); wt[j] = tp.y;                 // a0 | a1 << 16
        if (wt[j]) base = min(base, o[j]);
    }
    if (base == 0x7fffffff) base = 0;
    uint32_t sel[4];
    bool bad = false;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int d = wt[j] ? o[j] - base : 0;
        bad |= d > 6;
        sel[j] = (uint32_t)d | ((uint32_t)(d + 1) << 4);
    }
    if (bad) atomicExch(a.err, 202);
    const int wb = base >> 2;
    const unsigned sh = 8u * (unsigned)(base & 3);
    for (int r = slot; r < nrows; r += SLOTS) {
        const uint32_t* Sw = (const uint32_t*)(S + r * scols) + wb;
        const uint32_t w0 = Sw[0], w1 = Sw[1], w2 = Sw[2];
        const uint32_t X0 = __funnelshift_r(w0, w1, sh), X1 = __funnelshift_r(w1, w2, sh);
        uint32_t hh[4];
#pragma unroll
        for (int j = 0; j < 4; j++) hh[j] = __dp2a_lo(wt[j], __byte_perm(X0, X1, sel[j]), 0u) >> 4;
        *(uint2*)(Hq + r * HQS + 4 * q) = make_uint2(hh[0] | (hh[1] << 16), hh[2] | (hh[3] << 16));
    }
}

// vertical pass (VResizeLinear 8u): ((b0*H0)>>16) + ((b1*H1)>>16) + 2 >> 2; (b*lo16(u))>>16 == umulhi(u<<16, b)
template <int TH, int NT = PYR_THREADS>
__device__ __forceinline__ void tile_vresize(const PyrArgs& a, const TapX* __restrict__ tapy, const unsigned short* __restrict__ Hq,
                                             uint8_t* __restrict__ T, int x0, int y0, int tid) {
    constexpr int TROWS = TH + 2 * HALO;
    constexpr int SLOTS = NT / (TBYTES / 4);
    const int q = tid % (TBYTES / 4), slot = tid / (TBYTES / 4);
    const bool live = slot < SLOTS && x0 - 4 + 4 * q < a.w + HALO;
    if (!live) return;
    const int tyEnd = min(TROWS, a.h + 2 * HALO - y0);             // T rows below the image + halo are never read
    for (int ty = slot; ty < tyEnd; ty += SLOTS) {
        const uint2 tp = *(const uint2*)(tapy + ty);
        const uint32_t b0 = tp.y & 0xffffu, b1 = tp.y >> 16;     // weights are in [0, 2048]
        const uint2 u0 = *(const uint2*)(Hq + (tp.x & 0xffffu) * HQS + 4 * q);
        const uint2 u1 = *(const uint2*)(Hq + (tp.x >> 16) * HQS + 4 * q);
        const uint32_t v0 = (__umulhi(u0.x << 16, b0) + __umulhi(u1.x << 16, b1) + 2u) >> 2;
        const uint32_t v1 = (__umulhi(u0.x & 0xffff0000u, b0) + __umulhi(u1.x & 0xffff0000u, b1) + 2u) >> 2;
        const uint32_t v2 = (__umulhi(u0.y << 16, b0) + __umulhi(u1.y << 16, b1) + 2u) >> 2;
        const uint32_t v3 = (__umulhi(u0.y & 0xffff0000u, b0) + __umulhi(u1.y & 0xffff0000u, b1) + 2u) >> 2;
        *(uint32_t*)(T + ty * TS + 4 * q) = v0 | (v1 << 8) | (v2 << 16) | (v3 << 24);
    }
}

// shared memory carve-up (bytes).  T is live from the resize to the row pass; everything else lives in one region R
// behind it that is used twice: [tap tables | Hq | staged source rows] while T is being built, then
// [Hs | pair planes PA, PB] once T is complete.  54 KB per CTA at every level -> four CTAs per SM.
__host__ __device__ inline size_t pyr_smem_T(int trows) { return (size_t)trows * TS; }
__host__ __device__ inline size_t pyr_smem_taps() { return (size_t)TBYTES * sizeof(TapX) + 40 * sizeof(TapX); }
__host__ __device__ inline size_t pyr_smem_Hs(int trows) { return (size_t)trows * TW * 2; }
__host__ __device__ inline size_t pyr_smem_Hq(int srcRows) { return (((size_t)srcRows * HQS * 2) + 15) & ~(size_t)15; }
__host__ __device__ inline size_t pyr_smem_planes(int trows) { return (size_t)2 * trows * 32 * sizeof(uint4); }     // PA + PB
__host__ __device__ inline size_t pyr_smem_bytes(int th, bool level0, int srcRows, int srcCols) {
    const int trows = th + 2 * HALO;
    const size_t late = pyr_smem_Hs(trows) + pyr_smem_planes(trows);
    if (level0) return pyr_smem_T(trows) + late;
    const size_t stage = (size_t)srcRows * srcCols + 16;            // +16: S[o0 + 1] of the last column
    const size_t early = pyr_smem_taps() + pyr_smem_Hq(srcRows) + stage;
    return pyr_smem_T(trows) + (early > late ? early : late);
}

// NT threads per CTA: 256 in batches; calls of a frame or two run the same phases with 512 (one tile row per warp in the FAST and
// blur passes, 15 row slots in the resize passes): the eight dependent launches are the call's latency, and a CTA's serial chain
// through its phases is most of a launch
template <bool LEVEL0, int TH, int NT = PYR_THREADS>
__global__ void __launch_bounds__(NT) pyr_level_kernel(PyrArgs a) {
    constexpr int TROWS = TH + 2 * HALO;
    constexpr int NW = NT / 32;
    static_assert(TH % NW == 0 && NT >= 192 + TROWS, "tile rows per warp / tap-table threads");
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t* T = smem;                                   // TROWS x TS level tile with halo
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
    const int frame = blockIdx.z;
    const int w = a.w, h = a.h;
    const uint8_t* src = a.src + (size_t)frame * a.srcFrameStride;
    unsigned short* Hs;                                  // TROWS x TW blur row sums, [group][A.lo A.hi B.lo B.hi]
    uint4* PA;                                           // TROWS x 32 pair planes (see fast_score4); PB = PA + TROWS * 32

    Hs = (unsigned short*)(smem + pyr_smem_T(TROWS));
    PA = (uint4*)(smem + pyr_smem_T(TROWS) + pyr_smem_Hs(TROWS));
    if (LEVEL0) {
        // level 0 is the caller's image: copy with REFLECT_101 halo (copyMakeBorder, :1372); byte b of a T row is
        // column x0 - 4 + b.  Rows arrive at any alignment (the device copy keeps the host row step): interior
        // bytes go as aligned word loads + funnel shifts, the image edges byte by byte through reflect101.
        auto edge_word = [&](const uint8_t* row, int xb) {
            uint32_t v = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int x = xb + j;
                if (x >= -HALO && x < w + HALO) v |= (uint32_t)__ldg(row + reflect101(x, w)) << (8 * j);
            }
            return v;
        };
        // pass A: bytes 0..127 of every T row, two rows per warp step (half-warp = row, lane = 8 bytes)
        {
            const int half = lane >> 4, c = lane & 15;
            const int xb = x0 - 4 + 8 * c;
            const bool inner = xb >= 4 && xb + 11 < w;       // the three aligned words stay inside the row
            constexpr int NJ = (TROWS + 2 * NW - 1) / (2 * NW);      // two rows per warp and step
            uint2 v[NJ];
#pragma unroll
            for (int j = 0; j < NJ; j++) {
                const int ty = 2 * (wid + NW * j) + half;
                const int y = y0 - HALO + ty;
                v[j] = make_uint2(0u, 0u);
                if (ty < TROWS && y < h + HALO) {
                    const uint8_t* row = src + (size_t)reflect101(y, h) * a.srcPitch + a.srcX0;
                    if (inner) {
                        const uint8_t* p = row + xb;
                        const unsigned mis = (unsigned)((size_t)p & 3);
                        const uint32_t* ap = (const uint32_t*)(p - mis);
                        const uint32_t w0 = __ldg(ap), w1 = __ldg(ap + 1), w2 = __ldg(ap + 2);
                        v[j] = make_uint2(__funnelshift_r(w0, w1, 8 * mis), __funnelshift_r(w1, w2, 8 * mis));
                    } else {
                        v[j] = make_uint2(edge_word(row, xb), edge_word(row, xb + 4));
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < NJ; j++) {
                const int ty = 2 * (wid + NW * j) + half;
                if (ty < TROWS) *(uint2*)(T + ty * TS + 8 * c) = v[j];
            }
        }
        // pass B: bytes 128..135 (the right halo) of every T row
        if (tid < 2 * TROWS) {
            const int ty = tid >> 1, q = 32 + (tid & 1);
            const int y = y0 - HALO + ty, xb = x0 - 4 + 4 * q;
            uint32_t v = 0;
            if (y < h + HALO) {
                const uint8_t* row = src + (size_t)reflect101(y, h) * a.srcPitch + a.srcX0;
                if (xb + 7 < w) {
                    const uint8_t* p = row + xb;
                    const unsigned mis = (unsigned)((size_t)p & 3);
                    const uint32_t* ap = (const uint32_t*)(p - mis);
                    v = __funnelshift_r(__ldg(ap), __ldg(ap + 1), 8 * mis);
                } else {
                    v = edge_word(row, xb);
                }
            }
            *(uint32_t*)(T + ty * TS + 4 * q) = v;
        }
    } else {
        TapX* tapx = (TapX*)(smem + pyr_smem_T(TROWS));
        TapX* tapy = tapx + TBYTES;
        unsigned short* Hq = (unsigned short*)(smem + pyr_smem_T(TROWS) + pyr_smem_taps());   // source rows after the horizontal pass (h >> 4)
        uint8_t* S = (uint8_t*)Hq + pyr_smem_Hq(a.srcTileRows);                          // staged source rows
        // source rows/columns this tile touches
        const int ylo = max(0, y0 - HALO), yhi = min(h - 1, y0 + TH + HALO - 1);
        const int xlo = max(0, x0 - HALO), xhi = min(w - 1, x0 + TW + HALO - 1);
        int sy_lo, sy_hi, sx_lo, sx_hi;
        if (a.haveTiles) {                                     // from the kernel arguments: no memory round trip before the fetch
            const uint32_t ex = a.tileX[blockIdx.x], ey = a.tileY[blockIdx.y];
            sx_lo = (int)(ex & 0xffffu); sx_hi = (int)(ex >> 16); sy_lo = (int)(ey & 0xffffu); sy_hi = (int)(ey >> 16);
        } else {
            sy_lo = a.ytab[ylo].s0; sy_hi = a.ytab[yhi].s1;
            sx_lo = a.xtab[xlo].s0 & ~15; sx_hi = a.xtab[xhi].s1;
        }
        const int scols = a.srcTileCols;                       // multiple of 16
        const int nvec = (sx_hi - sx_lo) / 16 + 1;
        const int nrows = sy_hi - sy_lo + 1;
        if (nrows > a.srcTileRows || nvec * 16 > scols) { if (tid == 0) atomicExch(a.err, 201); return; }   // block-uniform
        tile_build_taps<TH>(a, tapx, tapy, x0, y0, sx_lo, sy_lo, tid);     // constant tables: before the wait
        pdl_wait();                                             // level l - 1 is complete (few-frames schedule, pdl.cuh)
        pdl_launch_dependents();
        for (int i = tid; i < nrows * nvec; i += NT) {
            const int r = i / nvec, c = i - r * nvec;
            const uint4 v = __ldg((const uint4*)(src + (size_t)(sy_lo + r) * a.srcPitch + a.srcX0 + sx_lo + 16 * c));
            *(uint4*)(S + r * scols + 16 * c) = v;
        }
        __syncthreads();
        tile_hresize<NT>(a, S, scols, nrows, tapx, Hq, x0, tid);
        __syncthreads();
        tile_vresize<TH, NT>(a, tapy, Hq, T, x0, y0, tid);
    }
    __syncthreads();

    tile_store_plane<TH>(a, T, frame, x0, y0, tid);
    uint4* PB = PA + TROWS * 32;
    tile_row_pass<TH, NT>(T, Hs, PA, PB, h, y0, wid, lane);
    __syncthreads();      // planes and Hs complete
    tile_fast_pass<TH, NT>(a, PA, PB, frame, x0, y0, wid, lane);
    tile_vblur_pass<TH, NT>(a, Hs, frame, x0, y0, wid, lane);
}

// BORDER_REFLECT_101 frame of 19 pixels around every level (copyMakeBorder,
// :1364-1365,:1372-1373).  One warp per group of BORDER_ROWS consecutive bordered rows: the 19 top and
// bottom rows are full-width copies of the reflected interior row, every row gets its 2 x 19 side pixels.
// All loads of a group are issued before its stores (the kernel is latency-bound: a row is a few dozen bytes).
constexpr int BORDER_ROWS = 4;
struct BorderArgs {
    uint8_t* pyr;                 // frame 0 pyramid block
    unsigned long long pyrFrameStride;
    int nlevels;
    int grpOff[ORB_MAX_LEVELS + 1];    // prefix sum of row groups, ceil((h + 38) / BORDER_ROWS) per level
    unsigned long long pyrOff[ORB_MAX_LEVELS];
    int w[ORB_MAX_LEVELS], h[ORB_MAX_LEVELS], pitch[ORB_MAX_LEVELS];
};

__global__ void __launch_bounds__(256) border_kernel(BorderArgs a) {
    const int lane = threadIdx.x & 31;
    const int grp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (grp >= a.grpOff[a.nlevels]) return;
    int l = 0;
    while (grp >= a.grpOff[l + 1]) l++;
    const int by0 = (grp - a.grpOff[l]) * BORDER_ROWS;   // first row of the group in the bordered plane
    const int w = a.w[l], h = a.h[l], pitch = a.pitch[l];
    uint8_t* plane = a.pyr + (size_t)blockIdx.y * a.pyrFrameStride + a.pyrOff[l];
    const uint8_t* srow[BORDER_ROWS];
    uint8_t* drow[BORDER_ROWS];
    bool live[BORDER_ROWS], full[BORDER_ROWS];
#pragma unroll
    for (int k = 0; k < BORDER_ROWS; k++) {
        const int by = by0 + k;
        live[k] = by < h + 2 * ORB_EDGE;
        full[k] = live[k] && (by < ORB_EDGE || by >= ORB_EDGE + h);
        const int sy = live[k] ? reflect101(by - ORB_EDGE, h) : 0;
        srow[k] = plane + (size_t)(sy + ORB_EDGE) * pitch + ORB_XPAD;     // interior row it mirrors
        drow[k] = plane + (size_t)by * pitch + ORB_XPAD;
    }
    // top / bottom rows: the interior part as 16-byte vectors (rows start 16-byte aligned)
    const int nvec = (w + 15) / 16;
    bool any = false;
#pragma unroll
    for (int k = 0; k < BORDER_ROWS; k++) any |= full[k];
    if (any) {                                           // warp-uniform
        for (int i0 = 0; i0 < nvec; i0 += 32) {
            const int i = i0 + lane;
            uint4 v[BORDER_ROWS];
#pragma unroll
            for (int k = 0; k < BORDER_ROWS; k++) if (full[k] && i < nvec) v[k] = __ldcg((const uint4*)srow[k] + i);
#pragma unroll
            for (int k = 0; k < BORDER_ROWS; k++) if (full[k] && i < nvec) *((uint4*)drow[k] + i) = v[k];   // up to 15 bytes into the right strip: rewritten below
        }
        __syncwarp();
    }
    if (lane < ORB_EDGE) {
        uint8_t vl[BORDER_ROWS], vr[BORDER_ROWS];
#pragma unroll
        for (int k = 0; k < BORDER_ROWS; k++) if (live[k]) { vl[k] = srow[k][1 + lane]; vr[k] = srow[k][w - 2 - lane]; }
#pragma unroll
        for (int k = 0; k < BORDER_ROWS; k++) if (live[k]) {
            drow[k][-1 - lane] = vl[k];                                   // column -1-k mirrors column 1+k
            drow[k][w + lane] = vr[k];                                    // column w+k mirrors column w-2-k
        }
    }
}

}  // namespace orbx
