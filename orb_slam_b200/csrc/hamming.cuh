// Brute-force Hamming 2-NN: ORBmatcher::DescriptorDistance
// (reference src/ORBmatcher.cc:1728-1744: 8 x xor + popcount over 32-bit words)
// applied exhaustively, with the best / second-best scan of :205-226.
//
// Layout: every thread keeps 4 query descriptors (32 registers) and their
// top-2 state; the database streams through shared memory in 128-row tiles
// (cp.async, double buffered) and every warp reads a row as two broadcast
// 128-bit loads.  The inner loop is 8 LOP3.XOR + 8 LOP3 (four carry-save adders) + 4 POPC
// + 3 adds + 1 compare per pair: the bound is the ALU pipe (LOP3 issues at half rate) with
// the POPC pipe close behind, not HBM (a shard is read once per query tile).
// top-2 is kept as packed keys (distance << 23 | local row) so the update is
// min/max only, and equal distances resolve to the lowest row = the
// reference's strict '<' scan in ascending index.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/orbm.h"

namespace orbm {

constexpr int BF_THREADS = 256;
constexpr int BF_QPT = 4;                     // queries per thread
constexpr int BF_QTILE = BF_THREADS * BF_QPT; // queries per block
constexpr int BF_ROWS = 128;                  // database rows per shared-memory tile (4 KB)
constexpr unsigned long long BF_NONE = (256ull << 32) | 0xffffffffull;

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

// 3-input logic with the truth table spelled out (one LOP3 each; immLut = f(0xF0, 0xCC, 0xAA))
__device__ __forceinline__ uint32_t lop_xor(uint32_t a, uint32_t b) { uint32_t r; asm("lop3.b32 %0, %1, %2, 0, 0x3c;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ uint32_t lop_xor3(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
__device__ __forceinline__ uint32_t lop_maj(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0xe8;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }

struct Top2Key { unsigned long long k1, k2; };   // (distance << 32 | global row), ascending

__global__ void __launch_bounds__(BF_THREADS) bf2nn_kernel(const uint32_t* __restrict__ q, int nq,
                                                           const uint32_t* __restrict__ db, unsigned long long ndb,
                                                           unsigned rowsPerSlice, uint32_t idxBase,
                                                           Top2Key* __restrict__ part) {
    __shared__ __align__(16) uint32_t tile[2][BF_ROWS * 8];
    const int tid = threadIdx.x;
    const unsigned slice = blockIdx.x;
    const int qbase = blockIdx.y * BF_QTILE;
    uint32_t qr[BF_QPT][8];
#pragma unroll
    for (int j = 0; j < BF_QPT; j++) {
        const int qi = qbase + j * BF_THREADS + tid;
        if (qi < nq) {
            const uint4 a = *(const uint4*)(q + (size_t)qi * 8), b = *(const uint4*)(q + (size_t)qi * 8 + 4);
            qr[j][0] = a.x; qr[j][1] = a.y; qr[j][2] = a.z; qr[j][3] = a.w;
            qr[j][4] = b.x; qr[j][5] = b.y; qr[j][6] = b.z; qr[j][7] = b.w;
        } else {
#pragma unroll
            for (int k = 0; k < 8; k++) qr[j][k] = 0;
        }
    }
    uint32_t b1[BF_QPT], b2[BF_QPT];
    uint32_t thr[BF_QPT];                              // distance of the current second best (512 = none yet)
#pragma unroll
    for (int j = 0; j < BF_QPT; j++) { b1[j] = 0xffffffffu; b2[j] = 0xffffffffu; thr[j] = 0x1ffu; }

    const unsigned long long r0 = (unsigned long long)slice * rowsPerSlice;
    const unsigned long long r1 = min(ndb, r0 + rowsPerSlice);
    const unsigned nrows = r1 > r0 ? (unsigned)(r1 - r0) : 0u;
    const unsigned ntiles = (nrows + BF_ROWS - 1) / BF_ROWS;
    const uint32_t* base = db + r0 * 8;

    auto issue = [&](unsigned t, int buf) {
        // 128 rows x 32 B = 256 x 16 B: one cp.async per thread; rows past the end are zero-filled
        const unsigned row = t * BF_ROWS + (tid >> 1);
        uint32_t* dst = &tile[buf][(tid >> 1) * 8 + (tid & 1) * 4];
        if (row < nrows) cp_async16(dst, base + (size_t)row * 8 + (tid & 1) * 4);
        else *(uint4*)dst = make_uint4(0, 0, 0, 0);
        cp_async_commit();
    };
    if (ntiles) issue(0, 0);
    for (unsigned t = 0; t < ntiles; t++) {
        const int buf = t & 1;
        if (t + 1 < ntiles) { issue(t + 1, buf ^ 1); cp_async_wait<1>(); } else cp_async_wait<0>();
        __syncthreads();
        const unsigned cnt = min((unsigned)BF_ROWS, nrows - t * BF_ROWS);
        const uint32_t rowBase = t * BF_ROWS;
#pragma unroll 2
        for (unsigned r = 0; r < cnt; r++) {
            const uint4 w0 = *(const uint4*)&tile[buf][r * 8], w1 = *(const uint4*)&tile[buf][r * 8 + 4];
            uint32_t d[BF_QPT];
            bool any = false;
#pragma unroll
            for (int j = 0; j < BF_QPT; j++) {
                // 256-bit Hamming distance with 4 POPC instead of 8: carry-save adders (LOP3 xor3 / majority) compress
                // the eight XOR words to two weight-1, one weight-2 and one weight-4 word first.  POPC costs four ALU
                // issue slots, a LOP3 one.
                // The adders are written as lop3.b32 with fixed truth tables (0x96 = a^b^c, 0xe8 = majority): left to
                // the compiler, the query XORs get folded into the adders as other 3-input functions and the pair ends
                // up with ~23 LOP3 instead of 8 XOR + 8 adder halves = 16 (profiles/r2_bf2nn_sass.md).
                const uint32_t x0 = lop_xor(qr[j][0], w0.x), x1 = lop_xor(qr[j][1], w0.y), x2 = lop_xor(qr[j][2], w0.z);
                const uint32_t x3 = lop_xor(qr[j][3], w0.w), x4 = lop_xor(qr[j][4], w1.x), x5 = lop_xor(qr[j][5], w1.y);
                const uint32_t x6 = lop_xor(qr[j][6], w1.z), x7 = lop_xor(qr[j][7], w1.w);
                const uint32_t s1 = lop_xor3(x0, x1, x2), c1 = lop_maj(x0, x1, x2);
                const uint32_t s2 = lop_xor3(x3, x4, x5), c2 = lop_maj(x3, x4, x5);
                const uint32_t s3 = lop_xor3(x6, x7, s1), c3 = lop_maj(x6, x7, s1);
                const uint32_t t2 = lop_xor3(c1, c2, c3), f4 = lop_maj(c1, c2, c3);
                d[j] = (__popc(s2) + __popc(s3)) + ((__popc(t2) + (__popc(f4) << 1)) << 1);
                any |= d[j] < thr[j];
            }
            // rows arrive in ascending index, so a row enters the top-2 only with a strictly smaller distance than the
            // current second best: after the first few hundred rows almost no warp takes this branch
            if (__any_sync(0xffffffffu, any)) {
#pragma unroll
                for (int j = 0; j < BF_QPT; j++) {
                    const uint32_t key = (d[j] << 23) | (rowBase + r);
                    b2[j] = min(b2[j], max(key, b1[j]));
                    b1[j] = min(b1[j], key);
                    thr[j] = b2[j] >> 23;
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < BF_QPT; j++) {
        const int qi = qbase + j * BF_THREADS + tid;
        if (qi < nq) {
            Top2Key o;
            const unsigned long long gb = (unsigned long long)idxBase + r0;
            o.k1 = b1[j] == 0xffffffffu ? BF_NONE : (((unsigned long long)(b1[j] >> 23)) << 32) | (gb + (b1[j] & 0x7fffffu));
            o.k2 = b2[j] == 0xffffffffu ? BF_NONE : (((unsigned long long)(b2[j] >> 23)) << 32) | (gb + (b2[j] & 0x7fffffu));
            part[(size_t)slice * nq + qi] = o;
        }
    }
}

// two smallest keys over all slices -> orbm_top2
__global__ void merge_slices_kernel(const Top2Key* __restrict__ part, int nslices, int nq, orbm_top2* __restrict__ out) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    unsigned long long m1 = BF_NONE, m2 = BF_NONE;
    for (int s = 0; s < nslices; s++) {
        const Top2Key p = part[(size_t)s * nq + qi];
        if (p.k1 < m1) { m2 = m1; m1 = p.k1; } else if (p.k1 < m2) m2 = p.k1;
        if (p.k2 < m1) { m2 = m1; m1 = p.k2; } else if (p.k2 < m2) m2 = p.k2;
    }
    orbm_top2 o;
    o.i1 = (uint32_t)(m1 & 0xffffffffull);
    o.d1 = (uint16_t)(m1 >> 32);
    o.d2 = (uint16_t)(m2 >> 32);
    out[qi] = o;
}

// exact merge of per-shard orbm_top2 results: best = min (d1, i1); second = next smallest distance
__global__ void merge_parts_kernel(const orbm_top2* __restrict__ parts, int nparts, int nq, orbm_top2* __restrict__ out) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    unsigned long long m1 = BF_NONE;
    unsigned second = 256;
    for (int p = 0; p < nparts; p++) {
        const orbm_top2 r = parts[(size_t)p * nq + qi];
        const unsigned long long k = ((unsigned long long)r.d1 << 32) | r.i1;
        if (k < m1) { if (m1 != BF_NONE) second = min(second, (unsigned)(m1 >> 32)); m1 = k; }
        else second = min(second, (unsigned)r.d1);
        second = min(second, (unsigned)r.d2);
    }
    orbm_top2 o;
    o.i1 = (uint32_t)(m1 & 0xffffffffull);
    o.d1 = (uint16_t)(m1 >> 32);
    o.d2 = (uint16_t)second;
    out[qi] = o;
}

// integer-pipe microbenchmark: 8 independent popc chains per thread
__global__ void popc_peak_kernel(unsigned* out, int iters, unsigned seed) {
    unsigned a[8], acc[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { a[k] = seed * (threadIdx.x + 1) + k * 0x9e3779b9u; acc[k] = 0; }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) { acc[k] += __popc(a[k] ^ acc[k]); }
    }
    unsigned s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s += acc[k];
    if (s == 0xdeadbeefu) out[0] = s;
}

// ALU-pipe microbenchmark: 8 independent lop3 chains per thread (the instruction that binds bf2nn_kernel: 16 per pair)
__global__ void lop3_peak_kernel(unsigned* out, int iters, unsigned seed) {
    unsigned a[8], b[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { a[k] = seed * (threadIdx.x + 1) + k * 0x9e3779b9u; b[k] = a[k] * 0x85ebca6bu + 1u; }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            a[k] = lop_maj(a[k], b[k], a[(k + 1) & 7]);
            b[k] = lop_xor3(b[k], a[k], b[(k + 3) & 7]);
        }
    }
    unsigned s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s += a[k] ^ b[k];
    if (s == 0xdeadbeefu) out[0] = s;
}

// The instruction mix of one (query, row) pair of bf2nn_kernel and nothing else - 8 LOP3 xor + 4 carry-save adders (8 LOP3) +
// 4 POPC + the weighted sum - on register operands, four independent pairs per iteration like the kernel's four queries per
// thread: the speed of light of that mix on this GPU (no shared-memory reads, no top-2 bookkeeping, no tile loop).
__global__ void __launch_bounds__(BF_THREADS) mix_peak_kernel(unsigned* out, int iters, unsigned seed) {
    uint32_t q[4][8], w[8], acc[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < 8; k++) {
        w[k] = seed * (threadIdx.x + 1) + k * 0x9e3779b9u;
#pragma unroll
        for (int j = 0; j < 4; j++) q[j][k] = w[k] * (0x85ebca6bu + 2 * j) + j;
    }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t x0 = lop_xor(q[j][0], w[0]), x1 = lop_xor(q[j][1], w[1]), x2 = lop_xor(q[j][2], w[2]);
            const uint32_t x3 = lop_xor(q[j][3], w[3]), x4 = lop_xor(q[j][4], w[4]), x5 = lop_xor(q[j][5], w[5]);
            const uint32_t x6 = lop_xor(q[j][6], w[6]), x7 = lop_xor(q[j][7], w[7]);
            const uint32_t s1 = lop_xor3(x0, x1, x2), c1 = lop_maj(x0, x1, x2);
            const uint32_t s2 = lop_xor3(x3, x4, x5), c2 = lop_maj(x3, x4, x5);
            const uint32_t s3 = lop_xor3(x6, x7, s1), c3 = lop_maj(x6, x7, s1);
            const uint32_t t2 = lop_xor3(c1, c2, c3), f4 = lop_maj(c1, c2, c3);
            acc[j] += (__popc(s2) + __popc(s3)) + ((__popc(t2) + (__popc(f4) << 1)) << 1);
        }
#pragma unroll
        for (int k = 0; k < 8; k++) w[k] += acc[k & 3] + 0x9e3779b9u;      // the next "row" depends on nothing expensive
    }
    if ((acc[0] ^ acc[1] ^ acc[2] ^ acc[3]) == 0xdeadbeefu) out[0] = acc[0];
}

}  // namespace orbm
