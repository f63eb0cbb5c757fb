// Orientation + descriptor + output assembly, one warp per retained keypoint:
//   IC_Angle            reference src/ORBextractor.cc:84-128  (int32 moments over the 749-pixel disc, fastAtan2)
//   computeOrbDescriptor                    :141-192  (256 rotated BRIEF tests on the blurred level)
//   operator() tail                         :1013-1025, :1290-1333 (border offset, octave, size, pt * scale,
//                                           levels concatenated in ascending order)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "geom.h"
#include "pdl.cuh"
#include "../../include/orbx.h"

namespace orbx {

__constant__ signed char c_pattern[1024];   // bit_pattern_31_ as (x,y) pairs, :196-454
__constant__ int c_umax[ORB_HALF_PATCH + 1];   // :522-539

struct DescArgs {
    const uint8_t* pyr; unsigned long long pyrFrameStride;
    const uint8_t* blur; unsigned long long planeFrameStride;
    const uint32_t* sel; const int* selCount;
    orbx_keypoint* kps; uint8_t* desc; int* nOut;
    int cap, nlevels, selTotal;
    int selOff[ORB_MAX_LEVELS + 1], pitch[ORB_MAX_LEVELS];
    unsigned long long pyrOff[ORB_MAX_LEVELS], planeOff[ORB_MAX_LEVELS];
    float scale[ORB_MAX_LEVELS], patchSize[ORB_MAX_LEVELS];
};

// cv::fastAtan2 scalar path (OpenCV 4.13 core/mathfuncs_core atan_f32): every
// operation rounded to float32, no FMA contraction.
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float scale = (float)(180.0 / 3.1415926535897932384626433832795);
    const float p1 = __fmul_rn(0.9997878412794807f, scale), p3 = __fmul_rn(-0.3258083974640975f, scale);
    const float p5 = __fmul_rn(0.1555786518463281f, scale), p7 = __fmul_rn(-0.04432655554792128f, scale);
    const float eps = 2.2204460492503131e-16f;   // (float)DBL_EPSILON
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

constexpr int DP_ROWS = 2 * ORB_EDGE + 1;      // blurred patch: 39 rows, |offset| <= ceil(13*sqrt(2)) = 19
constexpr int DP_VECS = 4;                     // 39 bytes starting anywhere inside a 16-byte vector fit in 4 vectors
constexpr int DP_SV = 5;                       // staged row stride in vectors: 80 bytes spreads the rows over all banks
constexpr int DP_STRIDE = DP_SV * 16;
constexpr int IP_ROWS = 2 * ORB_HALF_PATCH + 1;   // intensity-centroid disc: 31 rows
constexpr int IP_VECS = 3;                     // 31 bytes starting anywhere inside a 16-byte vector fit in 3 vectors
// keypoint slots per warp (template parameter of describe_kernel): 4 or 8 - chosen by the host from the size of the call

__device__ uint32_t g_pattern_s8[256];         // bit_pattern_31_ as [pair k][descriptor byte] -> bytes (x0, y0, x1, y1) + 16 of test 8*byte + k
// IC_Angle weights of disc row v = r - 15 (r = 0..30; row 31 is all zero), byte i <-> u = i - 15:
// [r][0..1] = 1 inside the disc (|u| <= umax[|v|]), [r][2..3] = u + 15 inside the disc
__device__ uint4 g_icw[32][4];

// byte n of w holds a pattern coordinate + 16 (1..31): dropped into the low mantissa byte of 1.5 * 2^23 it reads as the
// float 12582912 + (coordinate + 16), exactly; one subtraction leaves the coordinate as a float.  PRMT + FADD instead of
// PRMT + I2F: the conversion unit runs at a quarter of the FP32 rate and was 32 conversions per keypoint and lane.
template <unsigned SEL>
__device__ __forceinline__ float coord_f(uint32_t w) {
    uint32_t r;
    asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(w), "r"(0x4B400000u), "n"(SEL));   // volatile: keeps the 32 unpacked values out of loop-invariant registers
    return __fsub_rn(__uint_as_float(r), 12582928.0f);
}

// rowsum = sum of the row's disc pixels, su = sum of (u + 15) * pixel: 16 DP4A over the 32 bytes that start `sh/8` bytes
// into word WQ of the 12 loaded words
template <int WQ>
__device__ __forceinline__ void ic_row(const uint32_t (&W)[12], unsigned sh, const uint4 (&wt)[4], uint32_t& rowsum, uint32_t& su) {
    const uint32_t one[8] = {wt[0].x, wt[0].y, wt[0].z, wt[0].w, wt[1].x, wt[1].y, wt[1].z, wt[1].w};
    const uint32_t uw[8] = {wt[2].x, wt[2].y, wt[2].z, wt[2].w, wt[3].x, wt[3].y, wt[3].z, wt[3].w};
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const uint32_t d = __funnelshift_r(W[WQ + k], W[WQ + k + 1], sh);
        rowsum = __dp4a(d, one[k], rowsum);
        su = __dp4a(d, uw[k], su);
    }
}

template <int KPW>
__global__ void __launch_bounds__(256, 4) describe_kernel(DescArgs a) {
    // A warp owns KPW consecutive selection slots, slot `it` is lane `it`'s: its selection word, output index, moments, angle.
    //   phase A, slot by slot: the intensity-centroid disc is staged in shared memory with aligned, coalesced 128-bit
    //            loads so that every lane picks up one disc row as three vectors; the moments go to lane `it`;
    //   then lanes 0..KPW-1 run fastAtan2 and the sin/cos of their own slot side by side (the only long scalar chain
    //            of a keypoint: in a batch it costs a warp one pass for 32 keypoints instead of one per keypoint);
    //   phase B, slot by slot: the blurred patch is staged the same way for the 512 rotated-BRIEF gathers.
    __shared__ uint4 s_icw[32 * 4];
    __shared__ uint4 s_blur[8][DP_ROWS * DP_SV];
    __shared__ uint4 s_img[8][32 * IP_VECS];
    // transposed to [vector j][row]: the lanes of a warp read consecutive 16-byte vectors (row-major, 64 bytes per row, the eight
    // lanes of a quarter-warp hit two banks groups four times each: 16 wavefronts per read instead of 4)
    if (threadIdx.x < 128) s_icw[(threadIdx.x & 3) * 32 + (threadIdx.x >> 2)] = ((const uint4*)g_icw)[threadIdx.x];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int frame = blockIdx.y;
    pdl_wait();                                          // the selections are complete (few-frames schedule, pdl.cuh)
    // level counts: lane k fetches count k
    const int myCnt = lane < a.nlevels ? a.selCount[frame * a.nlevels + lane] : 0;
    if (blockIdx.x == 0 && wid == 0) {
        int all = myCnt;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) all += __shfl_xor_sync(0xffffffffu, all, d);
        if (lane == 0) a.nOut[frame] = all;
    }
    if (lane >= 31) {                                    // row 31 of the disc rows does not exist: zero it once
        s_img[wid][31 * IP_VECS] = make_uint4(0, 0, 0, 0); s_img[wid][31 * IP_VECS + 1] = make_uint4(0, 0, 0, 0);
        s_img[wid][31 * IP_VECS + 2] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    const int slot0 = (blockIdx.x * 8 + wid) * KPW;
    if (slot0 >= a.selTotal) return;                     // warp-uniform, after the only block-wide barrier
    // ---- this lane's slot: level, index inside the level, output row
    int myL = 0, myOut = -1;
    uint32_t myC = 0;
    {
        const int slot = slot0 + lane;
        if (lane < KPW && slot < a.selTotal) {
            while (slot >= a.selOff[myL + 1]) myL++;
            myC = a.sel[(size_t)frame * a.selTotal + slot];
        }
        const int i = slot - a.selOff[myL];
        int before = 0;                                  // keypoints of the lower levels: sum of their counts
#pragma unroll
        for (int l = 0; l < ORB_MAX_LEVELS; l++) {
            const int cl = __shfl_sync(0xffffffffu, myCnt, l);
            if (l < myL) before += cl;
        }
        const int cntL = __shfl_sync(0xffffffffu, myCnt, myL);
        if (lane < KPW && slot < a.selTotal && i < cntL && i + before < a.cap) myOut = i + before;   // (host reports ORBX_E_CAPACITY from nOut)
    }
    const unsigned live = __ballot_sync(0xffffffffu, myOut >= 0);
    if (!live) return;
    const int myX = (int)(myC & 0xfff) + ORB_BORDER_MIN, myY = (int)((myC >> 12) & 0xfff) + ORB_BORDER_MIN;   // :1021-1022

    // this lane's disc-row weights (the same for every keypoint): in registers for the whole warp's slots
    uint4 icw[4];
#pragma unroll
    for (int j = 0; j < 4; j++) icw[j] = s_icw[32 * j + lane];
    // ---- phase A: IC_Angle moments (:84-128): lane = disc row v + 15; m01 = sum v * rowsum(v), m10 = sum over rows of sum u * pixel
    int myM10 = 0, myM01 = 0;
#pragma unroll 1
    for (int it = 0; it < KPW; it++) {
        if (!((live >> it) & 1u)) continue;              // warp-uniform
        const int x = __shfl_sync(0xffffffffu, myX, it), y = __shfl_sync(0xffffffffu, myY, it), l = __shfl_sync(0xffffffffu, myL, it);
        const int pitch = a.pitch[l];
        // rows start 16-byte aligned at interior column 0 and every row has >= 32 bytes behind column w, so whole vectors
        // around the patch can be read (keypoints sit >= 19 px inside the level)
        const int ix0 = (x - ORB_HALF_PATCH) & ~15;
        constexpr int NI = (IP_ROWS * IP_VECS + 31) / 32;
        uint4 vi[NI];
        const uint8_t* im = a.pyr + (size_t)frame * a.pyrFrameStride + a.pyrOff[l] + (size_t)(ORB_EDGE + y - ORB_HALF_PATCH) * pitch + ORB_XPAD + ix0;
#pragma unroll
        for (int j = 0; j < NI; j++) {                   // lane t = row t/3, vector t%3
            const int t = lane + 32 * j, r = t / IP_VECS, v = t - r * IP_VECS;
            vi[j] = t < IP_ROWS * IP_VECS ? __ldg((const uint4*)(im + (size_t)r * pitch) + v) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int j = 0; j < NI; j++) { const int t = lane + 32 * j; if (t < IP_ROWS * IP_VECS) s_img[wid][t] = vi[j]; }
        __syncwarp();
        int m10, m01;
        {
            uint32_t W[12];
            const uint4 i0 = s_img[wid][lane * IP_VECS], i1 = s_img[wid][lane * IP_VECS + 1], i2 = s_img[wid][lane * IP_VECS + 2];
            W[0] = i0.x; W[1] = i0.y; W[2] = i0.z; W[3] = i0.w; W[4] = i1.x; W[5] = i1.y; W[6] = i1.z; W[7] = i1.w;
            W[8] = i2.x; W[9] = i2.y; W[10] = i2.z; W[11] = i2.w;
            const int off = (x - ORB_HALF_PATCH) - ix0;     // 0..15, warp-uniform
            const unsigned sh = 8u * (unsigned)(off & 3);

            uint32_t rowsum = 0, su = 0;
            switch (off >> 2) {
                case 0: ic_row<0>(W, sh, icw, rowsum, su); break;
                case 1: ic_row<1>(W, sh, icw, rowsum, su); break;
                case 2: ic_row<2>(W, sh, icw, rowsum, su); break;
                default: ic_row<3>(W, sh, icw, rowsum, su); break;
            }
            m10 = (int)su - ORB_HALF_PATCH * (int)rowsum;
            m01 = (lane - ORB_HALF_PATCH) * (int)rowsum;
        }
        __syncwarp();                                    // the staged disc may be overwritten by the next slot
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            m10 += __shfl_xor_sync(0xffffffffu, m10, d);
            m01 += __shfl_xor_sync(0xffffffffu, m01, d);
        }
        if (lane == it) { myM10 = m10; myM01 = m01; }
    }
    // ---- every live lane: orientation of its own keypoint and the rotation of the test pattern
    float myAngle = 0.f, myCa = 1.f, mySb = 0.f;
    if (myOut >= 0) {
        myAngle = fast_atan2_deg((float)myM01, (float)myM10);
        const float rad = __fmul_rn(myAngle, (float)(3.1415926535897932384626433832795 / 180.f));   // factorPI, :131
        double s_, c_;
        sincos((double)rad, &s_, &c_);                   // correctly-rounded stand-in for glibc cosf/sinf (:148)
        myCa = (float)c_; mySb = (float)s_;
    }
    __syncwarp();

    // ---- phase B: rotated BRIEF, lane i produces descriptor byte i (16 pattern points)
    const int brow = (lane >> 3) + 4 * ((lane >> 2) & 1);     // blurred-patch row (mod 8) this lane stages
    // test pair k of descriptor byte `lane` as four biased bytes (x0, y0, x1, y1) + 16, kept in registers for all slots
    uint32_t pat[8];
#pragma unroll
    for (int k = 0; k < 8; k++) pat[k] = g_pattern_s8[k * 32 + lane];
#pragma unroll 1
    for (int it = 0; it < KPW; it++) {
        if (!((live >> it) & 1u)) continue;              // warp-uniform
        const int x = __shfl_sync(0xffffffffu, myX, it), y = __shfl_sync(0xffffffffu, myY, it), l = __shfl_sync(0xffffffffu, myL, it);
        const int outIdx = __shfl_sync(0xffffffffu, myOut, it);
        const uint32_t c = __shfl_sync(0xffffffffu, myC, it);
        const float angle = __shfl_sync(0xffffffffu, myAngle, it);
        const float ca = __shfl_sync(0xffffffffu, myCa, it), sb = __shfl_sync(0xffffffffu, mySb, it);
        const int pitch = a.pitch[l];
        const int bx0 = (x - ORB_EDGE) & ~15;
        constexpr int NB = (DP_ROWS * DP_VECS + 31) / 32;
        uint4 vb[NB];
        {
            const uint8_t* bl = a.blur + (size_t)frame * a.planeFrameStride + a.planeOff[l] + (size_t)(y - ORB_EDGE) * pitch + ORB_XPAD + bx0;
            // lane -> (row, vector): a quarter-warp holds rows r and r + 4, whose staged vectors (80-byte row stride) fall
            // into disjoint banks, so the 128-bit shared stores below are conflict-free
            const uint8_t* p = bl + (size_t)brow * pitch + 16 * (lane & 3);
#pragma unroll
            for (int j = 0; j < NB; j++) {               // 8 rows per step
                vb[j] = (brow + 8 * j < DP_ROWS) ? __ldg((const uint4*)p) : make_uint4(0, 0, 0, 0);
                p += (size_t)8 * pitch;
            }
        }
#pragma unroll
        for (int j = 0; j < NB; j++) if (brow + 8 * j < DP_ROWS) s_blur[wid][(brow + 8 * j) * DP_SV + (lane & 3)] = vb[j];
        __syncwarp();
        // patch centre, minus the integer image of the rounding constant in both index terms
        const uint8_t* bl = (const uint8_t*)s_blur[wid] + ORB_EDGE * DP_STRIDE + (x - bx0) - (ptrdiff_t)0x4B400000 * (DP_STRIDE + 1);
        int val = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const float x0 = coord_f<0x7650>(pat[k]), y0 = coord_f<0x7651>(pat[k]);
            const float x1 = coord_f<0x7652>(pat[k]), y1 = coord_f<0x7653>(pat[k]);
            // cvRound(x*b + y*a), cvRound(x*a - y*b) (:166-168): adding 1.5 * 2^23 leaves the round-to-nearest-even integer
            // in the low mantissa bits (|value| < 2^22), so the conversion stays on the FP32 pipe
            const float MAGIC = 12582912.0f;
            const int r0 = __float_as_int(__fadd_rn(__fadd_rn(__fmul_rn(x0, sb), __fmul_rn(y0, ca)), MAGIC));
            const int q0 = __float_as_int(__fadd_rn(__fsub_rn(__fmul_rn(x0, ca), __fmul_rn(y0, sb)), MAGIC));
            const int r1 = __float_as_int(__fadd_rn(__fadd_rn(__fmul_rn(x1, sb), __fmul_rn(y1, ca)), MAGIC));
            const int q1 = __float_as_int(__fadd_rn(__fsub_rn(__fmul_rn(x1, ca), __fmul_rn(y1, sb)), MAGIC));
            const int t0 = bl[r0 * DP_STRIDE + q0], t1 = bl[r1 * DP_STRIDE + q1];
            val |= (t0 < t1) << k;
        }
        __syncwarp();                                    // the staged patch may be overwritten by the next slot
        const size_t o = (size_t)frame * a.cap + outIdx;
        a.desc[o * 32 + lane] = (uint8_t)val;
        if (lane < 7) {
            const float sc = a.scale[l];
            float f;
            int iv = 0;
            bool isInt = false;
            switch (lane) {
                case 0: f = l ? __fmul_rn((float)x, sc) : (float)x; break;     // keypoint->pt *= scale, :1322-1329
                case 1: f = l ? __fmul_rn((float)y, sc) : (float)y; break;
                case 2: f = a.patchSize[l]; break;
                case 3: f = angle; break;
                case 4: f = (float)(c >> 24); break;
                case 5: iv = l; isInt = true; f = 0; break;
                default: iv = -1; isInt = true; f = 0; break;
            }
            int* dst = (int*)(a.kps + o) + lane;
            *dst = isInt ? iv : __float_as_int(f);
        }
    }
}

}  // namespace orbx
