// Steps 0-2 of the quadtree stage (gather vToDistributeKeys, path codes, sort by path code; see quadtree_kernel.cuh) for
// LARGE levels, spread over a thread-block cluster of eight CTAs per (frame, level).
//
// quadtree_kernel runs one CTA per (frame, level); at 1080p / 4K the level-0 block holds 20 000 - 90 000 candidates and the
// single-CTA radix sort of their path codes in global memory is 55 - 80 % of the stage (phase clocks: 1.7 of 2.2 ms at 4K).
// The order-dependent part - replaying DistributeOctTree's list on the sorted ranges - stays with that one CTA; what it
// consumes is order-independent and is produced here:
//   0. every CTA scans the level's cell counts (cheap, redundant) and copies its eighth of the cells' survivor lists,
//   1. giving every candidate its path code and counting it into one of 16 384 buckets (global counters) on the way;
//   2a. bucket + rank sort (the codes are distinct): after a cluster barrier every CTA scans the bucket counts, places its
//      eighth of the keys behind a global fill counter per bucket and, after a second barrier, ranks its eighth of the placed
//      keys inside their buckets - one scattered write per key;
//   2b. inputs with a crowded bucket: LSD radix sort, 8 bits per pass: per-warp digit histograms of the CTA's contiguous key
//      range, the eight CTAs' digit totals exchanged through a small global scratch between two cluster barriers, then the
//      stable scatter.  Key i of the input lies in CTA floor(8i/M), warp by contiguous sub-range, so "digit, then CTA, then
//      warp, then position" is the input order: the sort is stable exactly like the single-CTA one.
// The cluster barrier (barrier.cluster, release/acquire) is the only inter-CTA synchronisation; the CTAs of a cluster are
// co-scheduled by the hardware, so the waits cannot deadlock.
#pragma once
#include <cooperative_groups.h>
#include "quadtree.cuh"

namespace orbx {

constexpr int QTP_CTAS = 8;                    // cluster size (portable maximum)
constexpr int QTP_THREADS = 512;
constexpr int QTP_WARPS = QTP_THREADS / 32;
constexpr int QTP_LOG = 14;                    // buckets of the bucket + rank sort (quadtree.cuh) per (frame, level): counters in global memory
constexpr int QTP_NB = 1 << QTP_LOG;

struct QtPresortArgs {
    QtArgs q;
    int levelBegin;                            // blockIdx.y = 0 is this level
    int* totals;                               // [level][frame][QTP_CTAS][256] digit totals of a radix pass
    int maxCells;                              // cells of the largest level of the launch (sizes the scan array)
    int* buckets;                              // [level][frame][2][QTP_NB], the launch's levels zeroed before it: bucket counts, fill counters
    int* which;                                // [frames][nlevels] out: bit 0 = buffer holding the sorted pairs, bit 1 = duplicates not checked yet
};

// the exclusive scan array serves the cell counts first and the bucket counts afterwards
__host__ __device__ inline int qtp_scan_words(int maxCells) { const int c = (maxCells + 31) & ~31; return c > QTP_NB + 32 ? c : QTP_NB + 32; }
__host__ __device__ inline size_t qtp_smem_bytes(int maxCells) {
    return (size_t)qtp_scan_words(maxCells) * 4 + (size_t)256 * QTP_WARPS * 4 + 256 * 4 * 2 + QTP_WARPS * 4 + 64;
}

// number of 8-bit passes the sort of a level takes (-> which ping-pong buffer holds the result)
__host__ __device__ inline int qtp_passes(int depth, int nIni) {
    int iniBits = 0;
    while ((1 << iniBits) < nIni) iniBits++;
    return (2 * depth + iniBits + 7) / 8;
}

__global__ void __launch_bounds__(QTP_THREADS) qt_presort_kernel(QtPresortArgs pa) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    const QtArgs& a = pa.q;
    extern __shared__ __align__(16) uint8_t psm[];
    int* offs = (int*)psm;                                            // exclusive scan of the level's cell counts, then of the bucket counts
    uint32_t* hist = (uint32_t*)(offs + qtp_scan_words(pa.maxCells)); // [256][QTP_WARPS]
    int* gbase = (int*)(hist + 256 * QTP_WARPS);                      // [256] global base of a digit for this CTA
    int* dsum = gbase + 256;                                          // [256] scratch of the digit scan
    int* warpTmp = dsum + 256;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int rank = (int)cluster.block_rank();
    const int frame = blockIdx.x / QTP_CTAS, l = pa.levelBegin + blockIdx.y;
    const int depth = a.depth[l], nIni = a.nIni[l];
    const size_t off = (size_t)frame * a.candFrameStride + a.candOff[l];
    uint32_t* cand = a.cand + off;
    uint32_t* kA = a.keyA + off; uint32_t* kB = a.keyB + off;
    uint32_t* vA = a.valA + off; uint32_t* vB = a.valB + off;
    // scratch of this (level, frame) pair: indexed by the level itself, so that two launches of one call (the few-frames schedule
    // runs level 0 on a side stream beside the other levels) never share a slot
    const size_t pairIdx = (size_t)l * (gridDim.x / QTP_CTAS) + frame;
    int* totals = pa.totals + pairIdx * QTP_CTAS * 256;
    int* gcnt = pa.buckets + pairIdx * 2 * QTP_NB;
    int* gfill = gcnt + QTP_NB;
    int iniBits = 0;
    while ((1 << iniBits) < nIni) iniBits++;
    const int bits = 2 * depth + iniBits;
    const int sh = max(0, bits - QTP_LOG);

    // ---- 0. gather: the cells' survivor lists in cell row-major order (:952-1002)
    const int nc = a.cellOff[l + 1] - a.cellOff[l];
    const uint32_t* cnt = a.cellCount + (size_t)frame * a.totalCells + a.cellOff[l];
    for (int c = tid; c < nc; c += QTP_THREADS) offs[c] = (int)cnt[c];
    __syncthreads();
    const int M = block_scan_excl<QTP_THREADS>(offs, nc, warpTmp);
    if (M > a.candCap[l]) {                              // (the same on every CTA of the cluster: all leave)
        if (tid == 0 && rank == 0) { atomicExch(a.err, 105); a.candCount[frame * a.nlevels + l] = 0; }
        return;
    }
    if (tid == 0 && rank == 0) a.candCount[frame * a.nlevels + l] = M;
    if (M <= 0) return;
    {
        const uint32_t* slots = a.slots + (size_t)frame * a.slotFrameStride + a.slotOff[l];
        const int scap = a.slotCap[l];
        const int nchunks = (nc + 31) >> 5;
        const int ch0 = nchunks * rank / QTP_CTAS, ch1 = nchunks * (rank + 1) / QTP_CTAS;
        // chunks of 32 cells (one per lane): inside a chunk the warp copies the concatenated lists element by element (lane t of a
        // step finds its cell by a shuffle search over the lanes' running totals): balanced however the survivors are spread
        for (int ch = ch0 + wid; ch < ch1; ch += QTP_WARPS) {
            const int c = ch * 32 + lane;
            const int nmine = c < nc ? (int)cnt[c] : 0;
            const int base = offs[ch * 32];
            int incl = nmine;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            const int T = __shfl_sync(0xffffffffu, incl, 31);
            for (int t0 = 0; t0 < T; t0 += 32) {
                const int t = t0 + lane;
                int j = 0;                               // first lane whose running total exceeds t
#pragma unroll
                for (int step = 16; step > 0; step >>= 1) {
                    const int v = __shfl_sync(0xffffffffu, incl, j + step - 1);
                    if (v <= t) j += step;
                }
                j = min(j, 31);
                const int startj = __shfl_sync(0xffffffffu, incl - nmine, j);
                if (t < T) {
                    // ---- 1. path code (sort key), (response, candidate index) sort value and the bucket count, in the same sweep
                    const uint32_t cw = slots[(size_t)(ch * 32 + j) * scap + (t - startj)];
                    const int i = base + t;
                    cand[i] = cw;
                    const uint32_t code = path_code_lookup(a, l, cw);
                    kA[i] = code;
                    vA[i] = (cw & 0xff000000u) | (uint32_t)i;
                    atomicAdd(&gcnt[code >> sh], 1);
                }
            }
        }
    }
    __threadfence();
    cluster.sync();                                      // vToDistributeKeys, codes and bucket counts complete

    const int i0 = (int)((long long)M * rank / QTP_CTAS), i1 = (int)((long long)M * (rank + 1) / QTP_CTAS);
    // ---- 2a. bucket + rank sort (quadtree.cuh): every CTA scans the bucket counts (cheap, redundant), places its eighth of the
    //      keys behind a global fill counter per bucket, and after a cluster barrier ranks its eighth of the placed keys inside
    //      their buckets.  One scattered write per key instead of the radix sort's three: the scattered writes of all the
    //      clusters are what the L2 limits here.
    int* bst = offs;
    int big = 0;
    for (int i = tid; i < QTP_NB; i += QTP_THREADS) { const int v = __ldcg(&gcnt[i]); bst[i] = v; big |= v > a.bucketMax; }
    if (!__syncthreads_or(big)) {
        block_scan_excl<QTP_THREADS>(bst, QTP_NB, warpTmp);
        if (tid == 0) bst[QTP_NB] = M;
        __syncthreads();
        for (int i = i0 + tid; i < i1; i += QTP_THREADS) {
            const uint32_t k = __ldcg(&kA[i]);
            const int b = (int)(k >> sh);
            const int p = bst[b] + atomicAdd(&gfill[b], 1);
            kB[p] = k; vB[p] = __ldcg(&vA[i]);
        }
        __threadfence();
        cluster.sync();                                  // every key sits in its bucket
        bool dup = false;
        for (int p = i0 + tid; p < i1; p += QTP_THREADS) {
            const uint32_t k = __ldcg(&kB[p]);
            const int b = (int)(k >> sh);
            const int s0 = bst[b], e0 = bst[b + 1];
            int r = 0;
            for (int j = s0; j < e0; j++) {
                const uint32_t kj = __ldcg(&kB[j]);
                r += (kj < k) || (kj == k && j < p);
                dup |= kj == k && j != p;
            }
            kA[s0 + r] = k; vA[s0 + r] = __ldcg(&vB[p]);
        }
        if (dup) atomicExch(a.err, 101);                 // distinct pixels must have distinct codes
        if (tid == 0 && rank == 0) pa.which[frame * a.nlevels + l] = 0;
        return;
    }
    if (tid == 0 && rank == 0) pa.which[frame * a.nlevels + l] = (qtp_passes(depth, nIni) & 1) | 2;

    // ---- 2b. a crowded bucket somewhere: stable LSD radix sort, the CTA's keys being the index range [i0, i1) of the current
    //      input buffer
    const int per = ((((i1 - i0) + QTP_WARPS - 1) / QTP_WARPS) + 31) & ~31;
    const int wb = min(i0 + wid * per, i1), we = min(wb + per, i1);
    int which = 0;
    for (int shift = 0; shift < bits; shift += 8) {
        const uint32_t* ki = which ? kB : kA; const uint32_t* vi = which ? vB : vA;
        uint32_t* ko = which ? kA : kB; uint32_t* vo = which ? vA : vB;
        for (int i = tid; i < 256 * QTP_WARPS; i += QTP_THREADS) hist[i] = 0;
        __syncthreads();
        constexpr int RU = 4;
        for (int base = wb; base < we; base += 32 * RU) {
            uint32_t kk[RU];
#pragma unroll
            for (int u = 0; u < RU; u++) { const int i = base + 32 * u + lane; kk[u] = i < we ? ki[i] : 0u; }
#pragma unroll
            for (int u = 0; u < RU; u++) {
                const int i = base + 32 * u + lane;
                if (base + 32 * u >= we) break;          // warp-uniform
                const bool valid = i < we;
                const unsigned act = __ballot_sync(0xffffffffu, valid);
                if (valid) {
                    const uint32_t d = (kk[u] >> shift) & 255u;
                    const unsigned peers = __match_any_sync(act, d);
                    if (lane == __ffs(peers) - 1) hist[d * QTP_WARPS + wid] += __popc(peers);
                }
                __syncwarp();
            }
        }
        __syncthreads();
        // digit totals of this CTA -> the cluster; per-warp exclusive offsets inside a digit stay in hist
        if (tid < 256) {
            int run = 0;
            for (int w = 0; w < QTP_WARPS; w++) { const int v = (int)hist[tid * QTP_WARPS + w]; hist[tid * QTP_WARPS + w] = (uint32_t)run; run += v; }
            totals[rank * 256 + tid] = run;
        }
        __threadfence();
        cluster.sync();
        if (tid < 256) {
            int all = 0, before = 0;
            for (int c = 0; c < QTP_CTAS; c++) {
                const int v = __ldcg(&totals[c * 256 + tid]);
                all += v;
                if (c < rank) before += v;
            }
            dsum[tid] = all; gbase[tid] = before;
        }
        __syncthreads();
        block_scan_excl<QTP_THREADS>(dsum, 256, warpTmp);            // keys of the whole level with a smaller digit
        if (tid < 256) {
            const int g0 = dsum[tid] + gbase[tid];
            for (int w = 0; w < QTP_WARPS; w++) hist[tid * QTP_WARPS + w] += (uint32_t)g0;
        }
        __syncthreads();
        for (int base = wb; base < we; base += 32 * RU) {
            uint32_t kk[RU], vv[RU];
#pragma unroll
            for (int u = 0; u < RU; u++) { const int i = base + 32 * u + lane; const bool in = i < we; kk[u] = in ? ki[i] : 0u; vv[u] = in ? vi[i] : 0u; }
#pragma unroll
            for (int u = 0; u < RU; u++) {
                const int i = base + 32 * u + lane;
                if (base + 32 * u >= we) break;          // warp-uniform
                const bool valid = i < we;
                const unsigned act = __ballot_sync(0xffffffffu, valid);
                if (valid) {
                    const uint32_t k = kk[u], v = vv[u];
                    const uint32_t d = (k >> shift) & 255u;
                    const unsigned peers = __match_any_sync(act, d);
                    const uint32_t pos = hist[d * QTP_WARPS + wid] + __popc(peers & ((1u << lane) - 1));
                    ko[pos] = k; vo[pos] = v;
                    __syncwarp(act);
                    if (lane == __ffs(peers) - 1) hist[d * QTP_WARPS + wid] += __popc(peers);
                }
                __syncwarp();
            }
        }
        __threadfence();
        cluster.sync();                                  // the pass's output is complete; the totals may be overwritten
        which ^= 1;
    }
}


}  // namespace orbx
