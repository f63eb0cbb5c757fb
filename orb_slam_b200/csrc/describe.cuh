// Orientation + descriptor + output assembly, one warp per retained keypoint:
//   IC_Angle            reference src/ORBextractor.cc:84-128  (int32 moments over the 749-pixel disc, fastAtan2)
//   computeOrbDescriptor                    :141-192  (256 rotated BRIEF tests on the blurred level)
//   operator() tail                         :1013-1025, :1290-1333 (border offset, octave, size, pt * scale,
//                                           levels concatenated in ascending order)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "geom.h"
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
constexpr int DP_WORDS = 11;                   // 39 bytes at any alignment fit in 11 aligned words
constexpr int IP_ROWS = 2 * ORB_HALF_PATCH + 1;   // intensity-centroid patch: 31 rows
constexpr int IP_WORDS = 9;                    // 31 bytes at any alignment fit in 9 aligned words

__device__ float4 g_pattern4[256];             // bit_pattern_31_ as [pair k][descriptor byte] -> (x0, y0, x1, y1) of test 8*byte + k

__global__ void __launch_bounds__(256) describe_kernel(DescArgs a) {
    // Both patches of a keypoint are staged in shared memory with aligned, coalesced 32-bit loads (all in flight at
    // once); the 961 moment reads and 512 rotated-BRIEF gathers then hit shared memory instead of L1/L2.
    __shared__ float4 s_pat[256];
    __shared__ uint32_t s_blur[8][DP_ROWS * DP_WORDS];
    __shared__ uint32_t s_img[8][IP_ROWS * IP_WORDS];
    s_pat[threadIdx.x] = g_pattern4[threadIdx.x];        // blockDim.x == 256
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int slot = blockIdx.x * (blockDim.x >> 5) + wid;
    const int frame = blockIdx.y;
    if (slot >= a.selTotal) return;
    int l = 0;
    while (slot >= a.selOff[l + 1]) l++;
    const int i = slot - a.selOff[l];
    const int* cnt = a.selCount + frame * a.nlevels;
    if (slot == 0 && lane == 0) {
        int n = 0;
        for (int k = 0; k < a.nlevels; k++) n += cnt[k];
        a.nOut[frame] = n;
    }
    if (i >= cnt[l]) return;
    int outIdx = i;
    for (int k = 0; k < l; k++) outIdx += cnt[k];
    if (outIdx >= a.cap) return;                      // host reports ORBX_E_CAPACITY from nOut
    const uint32_t c = a.sel[(size_t)frame * a.selTotal + slot];
    const int x = (int)(c & 0xfff) + ORB_BORDER_MIN, y = (int)((c >> 12) & 0xfff) + ORB_BORDER_MIN;   // :1021-1022
    const int pitch = a.pitch[l];

    // ---- stage the patches (rows start 16-byte aligned at interior column 0: alignment = column & 3)
    const int bx0 = (x - ORB_EDGE) & ~3, ix0 = (x - ORB_HALF_PATCH) & ~3;     // both >= 0: keypoints sit >= 19 px inside
    {
        const uint8_t* bl = a.blur + (size_t)frame * a.planeFrameStride + a.planeOff[l] + (size_t)(y - ORB_EDGE) * pitch + ORB_XPAD + bx0;
        const uint8_t* im = a.pyr + (size_t)frame * a.pyrFrameStride + a.pyrOff[l] + (size_t)(ORB_EDGE + y - ORB_HALF_PATCH) * pitch + ORB_XPAD + ix0;
        constexpr int NB = (DP_ROWS * DP_WORDS + 31) / 32, NI = (IP_ROWS * IP_WORDS + 31) / 32;
        uint32_t vb[NB], vi[NI];
#pragma unroll
        for (int j = 0; j < NB; j++) {                   // all loads first: NB + NI independent requests in flight per lane
            const int t = lane + 32 * j, r = t / DP_WORDS, wv = t - r * DP_WORDS;
            vb[j] = t < DP_ROWS * DP_WORDS ? __ldg((const uint32_t*)(bl + (size_t)r * pitch) + wv) : 0u;
        }
#pragma unroll
        for (int j = 0; j < NI; j++) {
            const int t = lane + 32 * j, r = t / IP_WORDS, wv = t - r * IP_WORDS;
            vi[j] = t < IP_ROWS * IP_WORDS ? __ldg((const uint32_t*)(im + (size_t)r * pitch) + wv) : 0u;
        }
#pragma unroll
        for (int j = 0; j < NB; j++) { const int t = lane + 32 * j; if (t < DP_ROWS * DP_WORDS) s_blur[wid][t] = vb[j]; }
#pragma unroll
        for (int j = 0; j < NI; j++) { const int t = lane + 32 * j; if (t < IP_ROWS * IP_WORDS) s_img[wid][t] = vi[j]; }
    }
    __syncwarp();

    // ---- IC_Angle: lanes 0..30 own column u = lane-15 of the disc
    int m10 = 0, m01 = 0;
    const int u = lane - ORB_HALF_PATCH;
    if (lane < 31) {
        const uint8_t* img = (const uint8_t*)s_img[wid] + (x - ORB_HALF_PATCH - ix0) + lane;   // row 0 = v = -15
        const int au = abs(u);
        int colsum = 0;
#pragma unroll
        for (int v = -ORB_HALF_PATCH; v <= ORB_HALF_PATCH; v++) {
            if (au <= c_umax[abs(v)]) {
                const int val = img[(v + ORB_HALF_PATCH) * (IP_WORDS * 4)];
                colsum += val;
                m01 += v * val;
            }
        }
        m10 = u * colsum;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, d);
        m01 += __shfl_xor_sync(0xffffffffu, m01, d);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // ---- rotated BRIEF: lane i produces descriptor byte i (16 pattern points)
    float ca = 0.f, sb = 0.f;
    if (lane == 0) {
        const float rad = __fmul_rn(angle, (float)(3.1415926535897932384626433832795 / 180.f));   // factorPI, :131
        double s_, c_;
        sincos((double)rad, &s_, &c_);       // correctly-rounded stand-in for glibc cosf/sinf (:148)
        ca = (float)c_; sb = (float)s_;
    }
    ca = __shfl_sync(0xffffffffu, ca, 0);
    sb = __shfl_sync(0xffffffffu, sb, 0);
    const uint8_t* bl = (const uint8_t*)s_blur[wid] + ORB_EDGE * (DP_WORDS * 4) + (x - bx0);     // patch centre
    int val = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        // test pair k of byte `lane`: one 128-bit shared load (x0, y0, x1, y1)
        const float4 pw = s_pat[k * 32 + lane];
        const float x0 = pw.x, y0 = pw.y, x1 = pw.z, y1 = pw.w;
        const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, sb), __fmul_rn(y0, ca)));      // cvRound(x*b + y*a)
        const int q0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, ca), __fmul_rn(y0, sb)));      // cvRound(x*a - y*b)
        const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, sb), __fmul_rn(y1, ca)));
        const int q1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, ca), __fmul_rn(y1, sb)));
        const int t0 = bl[r0 * (DP_WORDS * 4) + q0], t1 = bl[r1 * (DP_WORDS * 4) + q1];
        val |= (t0 < t1) << k;
    }
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

}  // namespace orbx
