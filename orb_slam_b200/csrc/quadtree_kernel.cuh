// quadtree_kernel: see quadtree.cuh for the formulation.
#pragma once
#include "quadtree.cuh"
#include "pdl.cuh"
#include "quadtree_presort.cuh"

namespace orbx {

struct QtSmem {
    uint32_t* lo0; uint32_t* lo1; uint32_t* hi0; uint32_t* hi1; uint8_t* dep0; uint8_t* dep1;   // (no arrays: an index would push the struct into local memory)
    uint32_t* b1; uint32_t* b2; uint32_t* b3;
    int* sA; int* sB; int* sC;
    unsigned long long* skey;
    uint32_t* hist;
    int* warpTmp;
    int* misc;       // [0]=nToExpand [1]=minT [2]=ne counter
};

__host__ __device__ inline size_t qt_smem_bytes(int cap, int nthreads) {
    // lo,hi x2 (16B/node) + dep x2 (2B) + b1..b3 (12B) + sA..sC (12B) + skey (8B) = 50B/node
    return (size_t)cap * 52 + 256 * (nthreads / 32) * 4 + (nthreads / 32) * 4 + 64;
}

// children of node (lo,hi,dep): boundaries by digit; returns number of non-empty children, ne = those with >1 key
__device__ __forceinline__ int split_node(const uint32_t* keys, uint32_t lo, uint32_t hi, int dep, int depth,
                                          uint32_t& c1, uint32_t& c2, uint32_t& c3, int& ne) {
    const int shift = 2 * (depth - 1 - dep);
    c1 = digit_lower_bound(keys, lo, hi, shift, 1);
    c2 = digit_lower_bound(keys, c1, hi, shift, 2);
    c3 = digit_lower_bound(keys, c2, hi, shift, 3);
    const uint32_t n1 = c1 - lo, n2 = c2 - c1, n3 = c3 - c2, n4 = hi - c3;
    ne = (n1 > 1) + (n2 > 1) + (n3 > 1) + (n4 > 1);
    return (n1 > 0) + (n2 > 0) + (n3 > 0) + (n4 > 0);
}

// the same with the three searches side by side over [lo, hi): they are independent, so their chains of dependent reads overlap
__device__ __forceinline__ int split_node3(const uint32_t* keys, uint32_t lo, uint32_t hi, int dep, int depth,
                                           uint32_t& c1, uint32_t& c2, uint32_t& c3, int& ne) {
    const int shift = 2 * (depth - 1 - dep);
    uint32_t l1 = lo, h1 = hi, l2 = lo, h2 = hi, l3 = lo, h3 = hi;
    while ((l1 < h1) | (l2 < h2) | (l3 < h3)) {
        const uint32_t m1 = (l1 + h1) >> 1, m2 = (l2 + h2) >> 1, m3 = (l3 + h3) >> 1;
        const uint32_t d1 = l1 < h1 ? (keys[m1] >> shift) & 3u : 0u;
        const uint32_t d2 = l2 < h2 ? (keys[m2] >> shift) & 3u : 0u;
        const uint32_t d3 = l3 < h3 ? (keys[m3] >> shift) & 3u : 0u;
        if (l1 < h1) { if (d1 >= 1u) h1 = m1; else l1 = m1 + 1; }
        if (l2 < h2) { if (d2 >= 2u) h2 = m2; else l2 = m2 + 1; }
        if (l3 < h3) { if (d3 >= 3u) h3 = m3; else l3 = m3 + 1; }
    }
    c1 = l1; c2 = l2; c3 = l3;
    const uint32_t n1 = c1 - lo, n2 = c2 - c1, n3 = c3 - c2, n4 = hi - c3;
    ne = (n1 > 1) + (n2 > 1) + (n3 > 1) + (n4 > 1);
    return (n1 > 0) + (n2 > 0) + (n3 > 0) + (n4 > 0);
}

// writes the non-empty children of a node in push order n1..n4; push index p0.. ; list position = totalPushed-1-p
__device__ __forceinline__ void emit_children(uint32_t* nlo, uint32_t* nhi, uint8_t* ndep, uint32_t lo, uint32_t c1,
                                              uint32_t c2, uint32_t c3, uint32_t hi, int dep, int p0, int totalPushed) {
    const uint32_t bnd[5] = {lo, c1, c2, c3, hi};
    int p = p0;
#pragma unroll
    for (int q = 0; q < 4; q++)
        if (bnd[q + 1] > bnd[q]) {
            const int pos = totalPushed - 1 - p;
            nlo[pos] = bnd[q]; nhi[pos] = bnd[q + 1]; ndep[pos] = (uint8_t)(dep + 1);
            p++;
        }
}

#ifdef QT_PROFILE
__device__ long long g_qt_clock[64];
#define QT_MARK(k) do { if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) g_qt_clock[k] = clock64(); } while (0)
#else
#define QT_MARK(k) do { } while (0)
#endif

template <int NT>
__global__ void __launch_bounds__(NT) quadtree_kernel(QtArgs a) {
    constexpr int QT_THREADS = NT, QT_WARPS = NT / 32;
    QT_MARK(0);
    extern __shared__ __align__(16) uint8_t qsm[];
    const int tid = threadIdx.x;
    const int l = a.levelBegin + blockIdx.y, frame = blockIdx.x;   // level-major launch order: the long level-0 blocks start first
    const int cap = a.nodeCap;
    QtSmem s;
    {
        uint8_t* p = qsm;
        s.skey = (unsigned long long*)p; p += (size_t)cap * 8;
        s.lo0 = (uint32_t*)p; p += (size_t)cap * 4; s.lo1 = (uint32_t*)p; p += (size_t)cap * 4;
        s.hi0 = (uint32_t*)p; p += (size_t)cap * 4; s.hi1 = (uint32_t*)p; p += (size_t)cap * 4;
        s.b1 = (uint32_t*)p; p += (size_t)cap * 4; s.b2 = (uint32_t*)p; p += (size_t)cap * 4; s.b3 = (uint32_t*)p; p += (size_t)cap * 4;
        s.sA = (int*)p; p += (size_t)cap * 4; s.sB = (int*)p; p += (size_t)cap * 4; s.sC = (int*)p; p += (size_t)cap * 4;
        s.hist = (uint32_t*)p; p += 256 * QT_WARPS * 4;
        s.warpTmp = (int*)p; p += QT_WARPS * 4;
        s.misc = (int*)p; p += 64;
        s.dep0 = p; p += cap; s.dep1 = p;
    }
    pdl_wait();                                          // the cells' survivor lists are complete (few-frames schedule, pdl.cuh)
    pdl_launch_dependents();
    int* selCount = a.selCount + frame * a.nlevels + l;
    const int depth = a.depth[l], N = a.N[l], nIni = a.nIni[l];
    const size_t off = (size_t)frame * a.candFrameStride + a.candOff[l];
    uint32_t* cand = a.cand + off;
    uint32_t* kA = a.keyA + off; uint32_t* kB = a.keyB + off;
    uint32_t* vA = a.valA + off; uint32_t* vB = a.valB + off;

    // 0. vToDistributeKeys = the cells' survivor lists in cell row-major order (:952-1002).  Cells are taken in chunks
    //    of 32 (one per lane); chunk totals give every chunk its base, and inside a chunk the warp copies the
    //    concatenated lists element by element (lane t of a step finds its cell by a shuffle search over the lanes'
    //    running totals), so the copy is balanced however the survivors are spread over the cells.
    const bool pre = (a.presortMask >> l) & 1u;           // large level: gathered, coded and sorted by qt_presort_kernel
    int M, which;
    bool checkDup = true;
    if (pre) {
        M = a.candCount[frame * a.nlevels + l];
        QT_MARK(1);
        if (M <= 0) { if (tid == 0) *selCount = 0; return; }
        const int pw = a.preWhich[frame * a.nlevels + l];
        which = pw & 1;
        checkDup = (pw & 2) != 0;                         // (the bucket + rank sort has looked at every key's neighbours already)
    } else {
        const int nc = a.cellOff[l + 1] - a.cellOff[l];
        const uint32_t* cnt = a.cellCount + (size_t)frame * a.totalCells + a.cellOff[l];
        const int lane = tid & 31, wid = tid >> 5;
        const int nchunks = (nc + 31) >> 5;
        int* chunkSum = s.sA;                                    // nchunks <= cap is checked on the host
        for (int ch = wid; ch < nchunks; ch += QT_WARPS) {
            const int c = ch * 32 + lane;
            int v = c < nc ? (int)cnt[c] : 0;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
            if (lane == 0) chunkSum[ch] = v;
        }
        // bucket counters of the sort (step 2), filled while the candidates pass through registers
        int iniBits = 0;
        while ((1 << iniBits) < nIni) iniBits++;
        const int bits = 2 * depth + iniBits;
        const int sh = max(0, bits - QtBucket<NT>::LOG);
        int* bc = (int*)s.hist;
        for (int i = tid; i <= QtBucket<NT>::NB; i += QT_THREADS) bc[i] = 0;
        __syncthreads();
        int total = 0;
        for (int ch = 0; ch < nchunks; ch++) total += chunkSum[ch];
        M = total;
        if (M > a.candCap[l]) { if (tid == 0) { atomicExch(a.err, 105); *selCount = 0; a.candCount[frame * a.nlevels + l] = 0; } return; }
        if (tid == 0) a.candCount[frame * a.nlevels + l] = M;
        if (M <= 0) { if (tid == 0) *selCount = 0; return; }
        // sort arrays in shared memory when this launch reserved room for them (generic pointers from here on): the binary
        // searches of the passes below are chains of dependent loads, 10x shorter from shared memory than from L2
        if (M <= a.keyCap) {
            uint32_t* ks = (uint32_t*)(qsm + qt_smem_bytes(cap, NT));
            kA = ks; kB = ks + a.keyCap; vA = ks + 2 * a.keyCap; vB = ks + 3 * a.keyCap;
        }
        // 0 + 1. candidate list, path codes (sort keys) and the (response, candidate index) sort values in one sweep
        const uint32_t* slots = a.slots + (size_t)frame * a.slotFrameStride + a.slotOff[l];
        const int scap = a.slotCap[l];
        for (int ch = wid; ch < nchunks; ch += QT_WARPS) {
            int base = 0;
            for (int k = 0; k < ch; k++) base += chunkSum[k];
            const int c = ch * 32 + lane;
            const int nmine = c < nc ? (int)cnt[c] : 0;
            int incl = nmine;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            const int T = __shfl_sync(0xffffffffu, incl, 31);
            constexpr int GU = 4;                                // candidate words fetched ahead of the path-code arithmetic
            for (int t0 = 0; t0 < T; t0 += 32 * GU) {
                uint32_t cw[GU];
#pragma unroll
                for (int u = 0; u < GU; u++) {
                    cw[u] = 0;
                    if (t0 + 32 * u >= T) continue;              // warp-uniform
                    const int t = t0 + 32 * u + lane;
                    int j = 0;                                   // first lane whose running total exceeds t
#pragma unroll
                    for (int step = 16; step > 0; step >>= 1) {
                        const int v = __shfl_sync(0xffffffffu, incl, j + step - 1);
                        if (v <= t) j += step;
                    }
                    j = min(j, 31);
                    const int startj = __shfl_sync(0xffffffffu, incl - nmine, j);
                    if (t < T) cw[u] = slots[(size_t)(ch * 32 + j) * scap + (t - startj)];
                }
#pragma unroll
                for (int u = 0; u < GU; u++) {
                    const int t = t0 + 32 * u + lane;
                    if (t < T) {
                        const int i = base + t;
                        cand[i] = cw[u];
                        const uint32_t code = path_code_lookup(a, l, cw[u]);
                        kA[i] = code;
                        vA[i] = (cw[u] & 0xff000000u) | (uint32_t)i;
                        atomicAdd(&bc[code >> sh], 1);
                    }
                }
            }
        }
        __syncthreads();
        QT_MARK(1);
        QT_MARK(2);
        // 2. sort by code: buckets + ranks, or the radix sort when a bucket is large
        which = 0;
        if (!block_bucket_sort<NT>(kA, vA, kB, vB, M, sh, bc, s.warpTmp, a.bucketMax))
            which = block_radix_sort<NT>(kA, vA, kB, vB, M, bits, s.hist, s.warpTmp);
    }
    QT_MARK(3);
    const uint32_t* keys = which ? kB : kA;
    const uint32_t* vals = which ? vB : vA;
    // distinct pixels must have distinct codes (depth was sized for it on the host)
    if (checkDup)
        for (int i = tid + 1; i < M; i += QT_THREADS)
            if (keys[i] == keys[i - 1]) atomicExch(a.err, 101);

    // 3. initial nodes in push_back order, empty ones erased (:663-699)
    // (boundaries of the nIni ranges by nIni + 1 threads, then thread 0 compacts the non-empty ones in order)
    if (tid <= nIni && tid < cap) s.b1[tid] = (tid == nIni) ? (uint32_t)M : key_lower_bound(keys, M, (uint32_t)tid << (2 * depth));
    __syncthreads();
    if (tid == 0) {
        int n = 0;
        for (int i = 0; i < nIni; i++) {
            const uint32_t lo = s.b1[i], hi = s.b1[i + 1];
            if (hi > lo) { s.lo0[n] = lo; s.hi0[n] = hi; s.dep0[n] = 0; n++; }
        }
        s.misc[3] = n;
    }
    __syncthreads();
    int n = s.misc[3];
    int cur = 0;
    bool finish = false;
    int fresh = 0;            // number of nodes created by the last pass (front of the list)
    bool priority = false;
    const int SHORT = a.shortMax;                        // lists up to this long take the passes with register scans (one barrier per NT nodes)
    uint32_t* scr = s.hist;                              // block-scan scratch of those passes (the sort is done with it)
    int passNo = 0;
    QT_MARK(4);
#ifdef QT_PROFILE
    int qtIter = 0;
#endif

    while (!finish) {
#ifdef QT_PROFILE
        if (qtIter < 20) QT_MARK(8 + qtIter);
        if (tid == 0 && blockIdx.x == 0 && blockIdx.y == 0) g_qt_clock[40 + min(qtIter, 19)] = priority ? -n : n;
        qtIter++;
#endif
        uint32_t* clo = cur ? s.lo1 : s.lo0; uint32_t* chi = cur ? s.hi1 : s.hi0; uint8_t* cdep = cur ? s.dep1 : s.dep0;
        uint32_t* nlo = cur ? s.lo0 : s.lo1; uint32_t* nhi = cur ? s.hi0 : s.hi1; uint8_t* ndep = cur ? s.dep0 : s.dep1;
        const int prevSize = n;
        passNo ^= 1;
        if (tid == 0) { s.misc[0] = 0; s.misc[1] = 0x7fffffff; s.misc[2] = 0; s.misc[4 + passNo] = 0; }
        __syncthreads();
        if (!priority && n <= SHORT) {
            // ---- breadth pass of a short list (:726-789).  Node i belongs to thread i % NT (slice i / NT): the nodes' searches
            //      run side by side, and both prefix sums (children pushed / nodes that stay) are one packed register scan
            //      per slice with one barrier.  The passes of a level are a chain of block barriers; this form has three for a
            //      list of up to NT nodes.
            const int nsl = (n + NT - 1) / NT;
            int neLocal = 0;
            uint32_t base = 0;
            for (int sl = 0; sl < nsl; sl++) {
                const int i = sl * NT + tid;
                uint32_t pk = 0;
                if (i < n) {
                    const uint32_t lo = clo[i], hi = chi[i];
                    const int dp = cdep[i];
                    if (hi - lo > 1) {
                        if (dp >= depth) atomicExch(a.err, 102);
                        else {
                            uint32_t c1, c2, c3;
                            int ne;
                            pk = (uint32_t)split_node3(keys, lo, hi, dp, depth, c1, c2, c3, ne);
                            s.b1[i] = c1; s.b2[i] = c2; s.b3[i] = c3;
                            neLocal += ne;
                        }
                    } else pk = 1u << 16;
                }
                uint32_t tot;
                const uint32_t ex = base + block_excl_reg<NT>(pk, scr + 32 * (sl & 1), tot);
                base += tot;
                if (i < n) s.sB[i] = (int)ex;                          // low half: children pushed before mine, high half: stayers
            }
            neLocal = __reduce_add_sync(0xffffffffu, neLocal);
            if ((tid & 31) == 0 && neLocal) atomicAdd(&s.misc[4 + passNo], neLocal);
            const int totalPushed = (int)(base & 0xffffu), totalNoMore = (int)(base >> 16);
            if (totalPushed + totalNoMore > cap) { if (tid == 0) atomicExch(a.err, 103); break; }
            for (int sl = 0; sl < nsl; sl++) {
                const int i = sl * NT + tid;
                if (i < n) {
                    const uint32_t lo = clo[i], hi = chi[i], ex = (uint32_t)s.sB[i];
                    const int dp = cdep[i];
                    if (hi - lo > 1) {
                        if (dp < depth) emit_children(nlo, nhi, ndep, lo, s.b1[i], s.b2[i], s.b3[i], hi, dp, (int)(ex & 0xffffu), totalPushed);
                    } else { const int pos = totalPushed + (int)(ex >> 16); nlo[pos] = lo; nhi[pos] = hi; ndep[pos] = (uint8_t)dp; }
                }
            }
            __syncthreads();
            n = totalPushed + totalNoMore;
            fresh = totalPushed;
            cur ^= 1;
            const int nToExpand = s.misc[4 + passNo];                 // (this slot is cleared again two passes from now)
            if (n >= N || n == prevSize) finish = true;
            else if (n + nToExpand * 3 > N) priority = true;
        } else if (priority && n <= SHORT) {
            // ---- priority pass of a short list (:805-874).  The expandable nodes are put in walk order by ranking their
            //      (count, creation seq) keys against each other - the keys are distinct - instead of a sorting network; the
            //      walk's running list size, the split it stops at and the children's places all come from ONE scan of the
            //      children counts: after t + 1 splits the list has grown by (children so far) - (t + 1).
            const int nsl = (n + NT - 1) / NT;
            int cntLocal = 0;
            for (int sl = 0; sl < nsl; sl++) {
                const int i = sl * NT + tid;
                if (i < fresh) {
                    const uint32_t cnt = chi[i] - clo[i];
                    const unsigned long long key = cnt > 1 ? ((unsigned long long)cnt << 32) | (uint32_t)(fresh - 1 - i) : 0ull;
                    s.skey[i] = key;
                    cntLocal += key != 0;
                }
            }
            cntLocal = __reduce_add_sync(0xffffffffu, cntLocal);
            if ((tid & 31) == 0 && cntLocal) atomicAdd(&s.misc[2], cntLocal);
            __syncthreads();
            const int ne = s.misc[2];
            for (int sl = 0; sl < nsl; sl++) {
                const int i = sl * NT + tid;
                if (i < n) {
                    int rank = -1;
                    const unsigned long long key = i < fresh ? s.skey[i] : 0ull;
                    if (key) {
                        int r = 0;
                        for (int j = 0; j < fresh; j++) r += s.skey[j] > key;
                        rank = r;
                        s.sA[r] = i;                                  // walk order -> node
                    }
                    s.sC[i] = rank;                                   // node -> walk order
                }
            }
            __syncthreads();
            int* ncOf = (int*)s.skey;                                 // (the keys are not needed any more)
            const int nslE = (ne + NT - 1) / NT;
            uint32_t base = 0;
            for (int sl = 0; sl < nslE; sl++) {
                const int t = sl * NT + tid;
                uint32_t nc = 0;
                if (t < ne) {
                    const int pi = s.sA[t];
                    const int pdp = cdep[pi];
                    if (pdp >= depth) atomicExch(a.err, 102);
                    else {
                        uint32_t c1, c2, c3;
                        int dummy;
                        nc = (uint32_t)split_node3(keys, clo[pi], chi[pi], pdp, depth, c1, c2, c3, dummy);
                        s.b1[t] = c1; s.b2[t] = c2; s.b3[t] = c3;
                    }
                }
                uint32_t tot;
                const uint32_t ex = base + block_excl_reg<NT>(nc, scr + 32 * (sl & 1), tot);
                base += tot;
                if (t < ne) {
                    s.sB[t] = (int)ex; ncOf[t] = (int)nc;
                    if (n + (int)(ex + nc) - (t + 1) >= N) atomicMin(&s.misc[1], t);      // first split that fills the list (:867-868)
                }
            }
            __syncthreads();
            const int tStop = (s.misc[1] == 0x7fffffff) ? ne - 1 : s.misc[1];
            const int nProc = tStop + 1;                                 // 0 when ne == 0
            const int totalPushed = nProc > 0 ? s.sB[tStop] + ncOf[tStop] : 0;
            const int newSize = n + totalPushed - nProc;                 // the nProc split nodes leave the list
            if (newSize > cap) { if (tid == 0) atomicExch(a.err, 103); break; }
            base = 0;
            for (int sl = 0; sl < nsl; sl++) {
                const int i = sl * NT + tid;
                bool keep = false;
                if (i < n) { const int rank = s.sC[i]; keep = !(rank >= 0 && rank < nProc); }
                uint32_t tot;
                const uint32_t ex = base + block_excl_reg<NT>(keep ? 1u : 0u, scr + 64 + 32 * (sl & 1), tot);
                base += tot;
                if (keep) { const int pos = totalPushed + (int)ex; nlo[pos] = clo[i]; nhi[pos] = chi[i]; ndep[pos] = cdep[i]; }
            }
            for (int sl = 0; sl < nslE; sl++) {
                const int t = sl * NT + tid;
                if (t < nProc) {
                    const int pi = s.sA[t];
                    const int pdp = cdep[pi];
                    if (pdp < depth) emit_children(nlo, nhi, ndep, clo[pi], s.b1[t], s.b2[t], s.b3[t], chi[pi], pdp, s.sB[t], totalPushed);
                }
            }
            __syncthreads();
            n = newSize;
            fresh = totalPushed;
            cur ^= 1;
            if (n >= N || n == prevSize) finish = true;
        } else if (!priority) {
            // ---- breadth pass: split every node with more than one key (:726-789)
            int neLocal = 0;
            for (int i = tid; i < n; i += QT_THREADS) {
                const uint32_t lo = clo[i], hi = chi[i];
                int nc = 0, ne = 0;
                if (hi - lo > 1) {
                    if (cdep[i] >= depth) { atomicExch(a.err, 102); }
                    else {
                        uint32_t c1, c2, c3;
                        nc = split_node(keys, lo, hi, cdep[i], depth, c1, c2, c3, ne);
                        s.b1[i] = c1; s.b2[i] = c2; s.b3[i] = c3;
                    }
                    s.sA[i] = nc; s.sB[i] = 0;
                } else { s.sA[i] = 0; s.sB[i] = 1; }
                neLocal += ne;
            }
            if (neLocal) atomicAdd(&s.misc[0], neLocal);
            __syncthreads();
            const int totalPushed = block_scan_excl<NT>(s.sA, n, s.warpTmp);
            const int totalNoMore = block_scan_excl<NT>(s.sB, n, s.warpTmp);
            if (totalPushed + totalNoMore > cap) { if (tid == 0) atomicExch(a.err, 103); break; }
            for (int i = tid; i < n; i += QT_THREADS) {
                const uint32_t lo = clo[i], hi = chi[i];
                if (hi - lo > 1) emit_children(nlo, nhi, ndep, lo, s.b1[i], s.b2[i], s.b3[i], hi, cdep[i], s.sA[i], totalPushed);
                else { const int pos = totalPushed + s.sB[i]; nlo[pos] = lo; nhi[pos] = hi; ndep[pos] = cdep[i]; }
            }
            __syncthreads();
            n = totalPushed + totalNoMore;
            fresh = totalPushed;
            cur ^= 1;
            const int nToExpand = s.misc[0];
            if (n >= N || n == prevSize) finish = true;
            else if (n + nToExpand * 3 > N) priority = true;
            __syncthreads();
        } else {
            // ---- priority pass: split the recorded expandable nodes, largest first, newest first on ties,
            //      stopping as soon as the list holds N nodes (:805-874)
            for (int i = tid; i < fresh; i += QT_THREADS) {
                const uint32_t cnt = chi[i] - clo[i];
                if (cnt > 1) {
                    const int k = atomicAdd(&s.misc[2], 1);
                    s.skey[k] = ((unsigned long long)cnt << 32) | (uint32_t)(fresh - 1 - i);   // (count, creation seq)
                }
            }
            __syncthreads();
            const int ne = s.misc[2];
            int p2 = 1;
            while (p2 < ne) p2 <<= 1;
            for (int i = ne + tid; i < p2; i += QT_THREADS) s.skey[i] = 0;
            __syncthreads();
            // bitonic sort, descending.  Element i belongs to thread i % NT, so a compare-exchange distance j < 32 stays
            // inside one warp's 32-element blocks: those stages need only a warp barrier; a block barrier separates
            // them from the stages that cross warps.
            for (int k = 2; k <= p2; k <<= 1)
                for (int j = k >> 1; j > 0; j >>= 1) {
                    for (int i = tid; i < p2; i += QT_THREADS) {
                        const int ixj = i ^ j;
                        if (ixj > i) {
                            const unsigned long long x = s.skey[i], y = s.skey[ixj];
                            const bool desc = (i & k) == 0;
                            if (desc ? (x < y) : (x > y)) { s.skey[i] = y; s.skey[ixj] = x; }
                        }
                    }
                    const int jn = j > 1 ? (j >> 1) : k;             // distance of the next stage (k = first stage of the next round)
                    if (j >= 32 || (jn >= 32 && 2 * k <= p2) || (j == 1 && k == p2)) __syncthreads();
                    else __syncwarp();
                }
            for (int t = tid; t < ne; t += QT_THREADS) {
                const int i = fresh - 1 - (int)(s.skey[t] & 0xffffffffu);
                int nc = 0, dummy;
                if (cdep[i] >= depth) atomicExch(a.err, 102);
                else {
                    uint32_t c1, c2, c3;
                    nc = split_node(keys, clo[i], chi[i], cdep[i], depth, c1, c2, c3, dummy);
                    s.b1[t] = c1; s.b2[t] = c2; s.b3[t] = c3;
                }
                s.sA[t] = nc; s.sB[t] = nc - 1;
            }
            for (int i = tid; i < n; i += QT_THREADS) s.sC[i] = 1;     // keep flags over the old list
            __syncthreads();
            // inclusive running size after each split; first t reaching N ends the walk (:867-868)
            block_scan_excl<NT>(s.sB, ne, s.warpTmp);
            for (int t = tid; t < ne; t += QT_THREADS) {
                const int incl = s.sB[t] + (s.sA[t] - 1);
                if (n + incl >= N) atomicMin(&s.misc[1], t);
            }
            __syncthreads();
            const int tStop = (s.misc[1] == 0x7fffffff) ? ne - 1 : s.misc[1];
            const int nProc = tStop + 1;                                 // 0 when ne == 0
            int newSize = n;
            if (nProc > 0) newSize = n + s.sB[tStop] + (s.sA[tStop] - 1);
            __syncthreads();
            const int pushedAll = block_scan_excl<NT>(s.sA, ne, s.warpTmp);
            int totalPushed = 0;
            if (nProc > 0) totalPushed = (nProc == ne) ? pushedAll : s.sA[nProc];
            for (int t = tid; t < nProc; t += QT_THREADS) s.sC[fresh - 1 - (int)(s.skey[t] & 0xffffffffu)] = 0;
            __syncthreads();
            const int kept = block_scan_excl<NT>(s.sC, n, s.warpTmp);
            if (totalPushed + kept > cap) { if (tid == 0) atomicExch(a.err, 103); break; }
            for (int t = tid; t < nProc; t += QT_THREADS) {
                const int i = fresh - 1 - (int)(s.skey[t] & 0xffffffffu);
                emit_children(nlo, nhi, ndep, clo[i], s.b1[t], s.b2[t], s.b3[t], chi[i], cdep[i], s.sA[t], totalPushed);
            }
            for (int i = tid; i < n; i += QT_THREADS) {
                const bool keep = (i + 1 < n) ? (s.sC[i + 1] != s.sC[i]) : (kept != s.sC[i]);
                if (keep) { const int pos = totalPushed + s.sC[i]; nlo[pos] = clo[i]; nhi[pos] = chi[i]; ndep[pos] = cdep[i]; }
            }
            __syncthreads();
            n = newSize;
            fresh = totalPushed;
            cur ^= 1;
            if (n >= N || n == prevSize) finish = true;
        }
    }

    QT_MARK(5);
#ifdef QT_PROFILE
    if (tid == 0 && blockIdx.x == 0 && blockIdx.y == 0) { g_qt_clock[7] = qtIter; g_qt_clock[39] = M; }
#endif
    // 4. retain the best key of every node, list order (:885-904): highest response, earliest candidate on ties
    if (n > a.maxSel[l]) { if (tid == 0) atomicExch(a.err, 104); n = a.maxSel[l]; }
    uint32_t* sel = a.sel + (size_t)frame * a.selTotal + a.selOff[l];
    const uint32_t* flo = cur ? s.lo1 : s.lo0; const uint32_t* fhi = cur ? s.hi1 : s.hi0;
    for (int i = tid; i < n; i += QT_THREADS) {
        uint32_t best = 0;
        for (uint32_t k = flo[i]; k < fhi[i]; k++) {
            const uint32_t v = vals[k] ^ 0x00ffffffu;     // response high, inverted candidate index low
            best = max(best, v);
        }
        sel[i] = cand[(best ^ 0x00ffffffu) & 0x00ffffffu];
    }
    if (tid == 0) *selCount = n;
    QT_MARK(6);
}

}  // namespace orbx
