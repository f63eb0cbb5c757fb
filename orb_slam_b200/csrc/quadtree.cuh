// DistributeOctTree (reference src/ORBextractor.cc:644-907, DivideNode :569-629)
// as one thread block per (frame, level).
//
// B200 formulation.  The reference moves keypoints between std::list nodes; a
// node's membership, however, depends only on the keypoint's coordinates and on
// the (deterministic) ceil-halving of the initial boxes.  So every candidate
// gets a PATH CODE - initial node index followed by one 2-bit child digit per
// subdivision depth (n1=0 UL, n2=1 UR, n3=2 BL, n4=3 BR, decided by the very
// comparisons of :603-617; by table: the column digits depend on x alone, the row
// digits on y alone) - the candidates are sorted by that code once (the codes of
// distinct pixels are distinct: buckets of the top bits, then ranks inside a bucket;
// a stable radix sort for inputs with a crowded bucket), and from then on every
// node of the tree is a contiguous range of the sorted array: splitting a node is
// three binary searches, no data moves.  The list
// itself (front-to-back order, push_front of children, erase of parents, the
// breadth passes and the "largest node first" passes with their early exit) is
// replayed exactly on ranges, with block-wide scans replacing the sequential
// walk.  Ties between equal-size nodes follow the canonical rule of SURVEY
// §8(a) E4: newest node first (the reference orders them by heap address).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "geom.h"

namespace orbx {

// block sizes of quadtree_kernel<NT>: small blocks for batches (many independent (frame, level) blocks per SM), large ones
// when a call holds only a few frames and the level-0 block is the critical path
constexpr int QT_THREADS_BATCH = 128;
constexpr int QT_THREADS_LATENCY = 512;
constexpr int QT_KEYS_SMEM = 4096;         // keys per sort array kept in shared memory by the latency variant (64 KB)

struct QtArgs {
    uint32_t* cand;                  // packed x:12|y:12|score:8, [frames][candTotal]: vToDistributeKeys, built here
    unsigned long long candFrameStride;
    int* candCount;                  // [frames][nlevels] (written here)
    // per-cell survivor lists of cell_nms_kernel (cells.cuh), concatenated in cell row-major order
    const uint32_t* cellCount;       // [frames][totalCells]
    uint32_t* cellOffs;              // [frames][totalCells] scratch: exclusive scan of cellCount inside a level
    const uint32_t* slots;           // [frames][slotTotal]
    unsigned long long slotFrameStride;
    int totalCells;
    int cellOff[ORB_MAX_LEVELS + 1];
    unsigned long long slotOff[ORB_MAX_LEVELS];
    int slotCap[ORB_MAX_LEVELS];
    int candCap[ORB_MAX_LEVELS];
    uint32_t* keyA; uint32_t* keyB;  // sort ping-pong, same layout as cand
    uint32_t* valA; uint32_t* valB;
    uint32_t* sel;                   // [frames][selTotal] retained candidates, packed like cand
    int* selCount;                   // [frames][nlevels]
    int* err;                        // device error flag
    int nlevels, selTotal, nodeCap;
    int levelBegin;                  // blockIdx.y = 0 is this level (a launch may cover a level range)
    unsigned presortMask;            // bit l: steps 0-2 of level l were done by qt_presort_kernel (quadtree_presort.cuh)
    const int* preWhich;             // [frames][nlevels] left by qt_presort_kernel: bit 0 = buffer with the sorted pairs, bit 1 = check for duplicates
    int keyCap;                      // keys per sort array that fit in the dynamic shared memory behind the node arrays (0 = none)
    int shortMax;                    // lists up to this long take the register-scan passes (0: always the general passes)
    int bucketMax;                   // largest bucket the bucket + rank sort accepts (0: always the radix sort)
    unsigned long long candOff[ORB_MAX_LEVELS];
    // path codes by table: the column comparisons of DivideNode depend on x alone and the row comparisons on y alone, so a
    // candidate's code is codeTab[codeXOff[l] + x] | codeTab[codeYOff[l] + y] (tables built once per frame size, path_code_tables)
    const uint32_t* codeTab;
    int codeXOff[ORB_MAX_LEVELS], codeYOff[ORB_MAX_LEVELS];
    int N[ORB_MAX_LEVELS], nIni[ORB_MAX_LEVELS], qH[ORB_MAX_LEVELS], depth[ORB_MAX_LEVELS];
    int maxSel[ORB_MAX_LEVELS], selOff[ORB_MAX_LEVELS];
    float hX[ORB_MAX_LEVELS];
};

// exclusive scan of arr[0..n) in shared memory by the whole block; returns the total.  Every warp scans one contiguous
// segment 32 elements at a time (conflict-free, carry in a register); the warp totals are exchanged once.
template <int NT>
__device__ int block_scan_excl(int* arr, int n, int* warpTmp) {
    constexpr int QT_THREADS = NT, QT_WARPS = NT / 32;
    (void)QT_THREADS;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int seg = (((n + QT_WARPS - 1) / QT_WARPS) + 31) & ~31;
    const int b = min(wid * seg, n), e = min(b + seg, n);
    int carry = 0;
    for (int i0 = b; i0 < e; i0 += 32) {
        const int i = i0 + lane;
        const int v = i < e ? arr[i] : 0;
        int inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += t;
        }
        if (i < e) arr[i] = carry + inc - v;
        carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    __syncthreads();                       // protects warpTmp reuse across calls
    if (lane == 0) warpTmp[wid] = carry;
    __syncthreads();
    int base = 0, total = 0;
#pragma unroll
    for (int k = 0; k < QT_WARPS; k++) { const int v = warpTmp[k]; if (k < wid) base += v; total += v; }
    if (base) for (int i = b + lane; i < e; i += 32) arr[i] += base;
    __syncthreads();
    return total;
}

// first index in [lo,hi) whose 2-bit digit at `shift` is >= q (digits are non-decreasing inside a node)
__device__ __forceinline__ uint32_t digit_lower_bound(const uint32_t* keys, uint32_t lo, uint32_t hi, int shift, uint32_t q) {
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (((keys[mid] >> shift) & 3u) >= q) hi = mid; else lo = mid + 1;
    }
    return lo;
}
__device__ __forceinline__ uint32_t key_lower_bound(const uint32_t* keys, uint32_t n, uint32_t v) {
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (keys[mid] >= v) hi = mid; else lo = mid + 1;
    }
    return lo;
}

// path code of one candidate: replays DivideNode's box arithmetic from the initial node down.  (Reference form of the code; the
// kernels read it from the tables below, which are this function evaluated per column and per row.)
__host__ __device__ __forceinline__ uint32_t path_code(int x, int y, float hX, int nIni, int qH, int depth) {
#ifdef __CUDA_ARCH__
    int ini = (int)__fdiv_rn((float)x, hX);                    // vpIniNodes[kp.pt.x/hX], :682
    ini = min(ini, nIni - 1);
    int ulx = (int)__fmul_rn(hX, (float)ini), urx = (int)__fmul_rn(hX, (float)(ini + 1));   // :667-668
#else
    int ini = (int)((float)x / hX);                            // (host: IEEE float ops, the library is built with -ffp-contract=off)
    ini = ini < nIni - 1 ? ini : nIni - 1;
    int ulx = (int)(hX * (float)ini), urx = (int)(hX * (float)(ini + 1));
#endif
    int uly = 0, bry = qH;
    uint32_t code = (uint32_t)ini;
    for (int t = 0; t < depth; t++) {
        const int midX = ulx + ((urx - ulx + 1) >> 1);           // ceil((UR.x-UL.x)/2), :572
        const int midY = uly + ((bry - uly + 1) >> 1);
        uint32_t q;
        if (x < midX) { urx = midX; q = 0; } else { ulx = midX; q = 1; }
        if (y < midY) { bry = midY; } else { uly = midY; q |= 2; }
        code = (code << 2) | q;
    }
    return code;
}

// The column digits (bit 0 of every 2-bit digit) and the initial node index depend on x alone, the row digits (bit 1) on y
// alone: path_code(x, y) = path_code(x, 0) & XMASK | path_code(0, y) & YMASK.  Host: fills xs[0..nx) and ys[0..ny).
inline void path_code_tables(float hX, int nIni, int qH, int depth, int nx, int ny, uint32_t* xs, uint32_t* ys) {
    uint32_t ymask = 0;
    for (int t = 0; t < depth; t++) ymask |= 2u << (2 * t);
    for (int x = 0; x < nx; x++) xs[x] = path_code(x, 0, hX, nIni, qH, depth) & ~ymask;
    for (int y = 0; y < ny; y++) ys[y] = path_code(0, y, hX, nIni, qH, depth) & ymask;
}
__device__ __forceinline__ uint32_t path_code_lookup(const QtArgs& a, int l, uint32_t cw) {
    return __ldg(&a.codeTab[a.codeXOff[l] + (int)(cw & 0xfff)]) | __ldg(&a.codeTab[a.codeYOff[l] + (int)((cw >> 12) & 0xfff)]);
}

// stable LSD radix sort of (key,val) pairs by `bits` low bits of key; returns 0 if the result is in A, 1 if in B
template <int NT>
__device__ int block_radix_sort(uint32_t* kA, uint32_t* vA, uint32_t* kB, uint32_t* vB, int M, int bits, uint32_t* hist /*256*QT_WARPS*/, int* warpTmp) {
    constexpr int QT_THREADS = NT, QT_WARPS = NT / 32;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int per = (((M + QT_WARPS - 1) / QT_WARPS) + 31) & ~31;
    const int wb = min(wid * per, M), we = min(wb + per, M);
    int which = 0;
    for (int shift = 0; shift < bits; shift += 8) {
        const uint32_t* ki = which ? kB : kA; const uint32_t* vi = which ? vB : vA;
        uint32_t* ko = which ? kA : kB; uint32_t* vo = which ? vA : vB;
        for (int i = tid; i < 256 * QT_WARPS; i += QT_THREADS) hist[i] = 0;
        __syncthreads();
        // both loops fetch RU 32-key chunks ahead of processing them: the fetches are independent, the histogram
        // updates are not
        constexpr int RU = 4;
        for (int base = wb; base < we; base += 32 * RU) {
            uint32_t kk[RU];
#pragma unroll
            for (int u = 0; u < RU; u++) { const int i = base + 32 * u + lane; kk[u] = i < we ? ki[i] : 0u; }
#pragma unroll
            for (int u = 0; u < RU; u++) {
                const int i = base + 32 * u + lane;
                if (base + 32 * u >= we) break;                     // warp-uniform
                const bool valid = i < we;
                const unsigned act = __ballot_sync(0xffffffffu, valid);
                if (valid) {
                    const uint32_t d = (kk[u] >> shift) & 255u;
                    const unsigned peers = __match_any_sync(act, d);
                    if (lane == __ffs(peers) - 1) hist[d * QT_WARPS + wid] += __popc(peers);
                }
                __syncwarp();
            }
        }
        __syncthreads();
        block_scan_excl<NT>((int*)hist, 256 * QT_WARPS, warpTmp);
        for (int base = wb; base < we; base += 32 * RU) {
            uint32_t kk[RU], vv[RU];
#pragma unroll
            for (int u = 0; u < RU; u++) { const int i = base + 32 * u + lane; const bool in = i < we; kk[u] = in ? ki[i] : 0u; vv[u] = in ? vi[i] : 0u; }
#pragma unroll
            for (int u = 0; u < RU; u++) {
                const int i = base + 32 * u + lane;
                if (base + 32 * u >= we) break;                     // warp-uniform
                const bool valid = i < we;
                const unsigned act = __ballot_sync(0xffffffffu, valid);
                if (valid) {
                    const uint32_t k = kk[u], v = vv[u];
                    const uint32_t d = (k >> shift) & 255u;
                    const unsigned peers = __match_any_sync(act, d);
                    const uint32_t pos = hist[d * QT_WARPS + wid] + __popc(peers & ((1u << lane) - 1));
                    ko[pos] = k; vo[pos] = v;
                    __syncwarp(act);
                    if (lane == __ffs(peers) - 1) hist[d * QT_WARPS + wid] += __popc(peers);
                }
                __syncwarp();
            }
        }
        __syncthreads();
        which ^= 1;
    }
    return which;
}

// ---- sort of DISTINCT keys by bucket, then rank --------------------------------------------------------------------------
// The path codes of distinct pixels are distinct (quadtree_kernel checks it, error 101), so the sorted order does not depend on
// the algorithm.  A counting pass scatters the keys into buckets of their top bits (shared-memory atomics; the order inside a
// bucket is whatever the atomics made it), then every key finds its final place by counting the smaller keys of its own bucket.
// A bucket is a small box of the level and holds a few keys; the rank step is quadratic in the bucket size, so the caller keeps
// the radix sort for inputs with a large bucket.  Two passes over the keys and one scan of the bucket counts instead of the
// radix sort's three rounds of (histogram by match_any, scan of 256 x warps counters, scatter by match_any).
template <int NT> struct QtBucket {
    static constexpr int LOG = NT >= 512 ? 11 : 9;       // NB + 1 counters must fit in the 256 x warps words of the radix histogram
    static constexpr int NB = 1 << LOG;
    static constexpr int MAXB = 64;                      // larger bucket anywhere -> radix sort (QtArgs::bucketMax)
    static constexpr int SHORT = 8 * NT;                 // QtArgs::shortMax
};

// bc[b] = number of keys of bucket b = key >> sh (counted by the caller while it produced kA / vA), bc[NB] unused.
// Returns false (nothing moved) when some bucket exceeds maxBucket; otherwise the sorted pairs are back in kA / vA.
template <int NT>
__device__ bool block_bucket_sort(uint32_t* kA, uint32_t* vA, uint32_t* kB, uint32_t* vB, int M, int sh, int* bc, int* warpTmp, int maxBucket) {
    constexpr int NB = QtBucket<NT>::NB;
    const int tid = threadIdx.x;
    int big = 0;
    for (int i = tid; i < NB; i += NT) big |= bc[i] > maxBucket;
    if (__syncthreads_or(big)) return false;
    block_scan_excl<NT>(bc, NB, warpTmp);                // bc[b] = first position of bucket b
    for (int i = tid; i < M; i += NT) {
        const uint32_t k = kA[i];
        const int p = atomicAdd(&bc[k >> sh], 1);
        kB[p] = k; vB[p] = vA[i];
    }
    __syncthreads();                                     // now bc[b] = end of bucket b = start of bucket b + 1
    for (int p = tid; p < M; p += NT) {
        const uint32_t k = kB[p];
        const int b = (int)(k >> sh);
        const int s0 = b ? bc[b - 1] : 0, e0 = bc[b];
        int r = 0;
        for (int j = s0; j < e0; j++) { const uint32_t kj = kB[j]; r += (kj < k) || (kj == k && j < p); }
        kA[s0 + r] = k; vA[s0 + r] = vB[p];
    }
    __syncthreads();
    return true;
}

// exclusive block scan of one register value per thread (one barrier); wt = NT / 32 words of shared scratch that no thread
// touches again before the next block barrier
template <int NT>
__device__ __forceinline__ uint32_t block_excl_reg(uint32_t mine, uint32_t* wt, uint32_t& total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) wt[wid] = inc;
    __syncthreads();
    uint32_t base = 0;
    total = 0;
#pragma unroll
    for (int k = 0; k < NT / 32; k++) { const uint32_t v = wt[k]; if (k < wid) base += v; total += v; }
    return base + inc - mine;
}

}  // namespace orbx
