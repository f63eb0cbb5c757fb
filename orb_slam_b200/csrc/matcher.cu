// Host side of the B200 Hamming matcher / Frame grid: extern "C" entry points of include/orbm.h.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/orbm.h"
#include "hamming.cuh"
#include "window.cuh"
#include "bow.cuh"
#include "initbatch.cuh"
#include "vocab.cuh"

using namespace orbm;

#define CKM(expr)                                                     \
    do {                                                              \
        cudaError_t e_ = (expr);                                      \
        if (e_ != cudaSuccess) {                                      \
            std::fprintf(stderr, "orbm: %s failed: %s (%s:%d)\n", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return (int)e_;                                           \
        }                                                             \
    } while (0)

struct DevBuf {
    void* p = nullptr; size_t cap = 0;
    int reserve(size_t n) {
        if (n <= cap) return 0;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, n);
        if (e != cudaSuccess) return (int)e;
        cap = n;
        return 0;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
struct PinBuf {
    void* p = nullptr; size_t cap = 0;
    int reserve(size_t n) {
        if (n <= cap) return 0;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMallocHost(&p, n);
        if (e != cudaSuccess) return (int)e;
        cap = n;
        return 0;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

struct orbm_context {
    int device = 0;
    int numSMs = 148;
    cudaStream_t stream = nullptr;
    DevBuf part, q, out, w0, w1, w2, w3, w4, w5;
    PinBuf h0, h1, hf;
    size_t lastPairs = 0;                    // candidate pairs the previous window search produced
    std::mutex mtx;
};

extern "C" {

int orbm_create(orbm_t** out, int device) {
    if (!out) return ORBX_E_INVALID;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return ORBX_E_NO_DEVICE;
    cudaDeviceProp prop;
    CKM(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return ORBX_E_NO_DEVICE;
    CKM(cudaSetDevice(device));
    orbm_context* m = new orbm_context();
    m->device = device; m->numSMs = prop.multiProcessorCount;
    cudaError_t rc = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking);
    if (rc != cudaSuccess) { delete m; return (int)rc; }
    *out = m;
    return ORBX_OK;
}

void orbm_destroy(orbm_t* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    cudaStreamSynchronize(m->stream);
    for (DevBuf* b : {&m->part, &m->q, &m->out, &m->w0, &m->w1, &m->w2, &m->w3, &m->w4, &m->w5}) b->release();
    m->h0.release(); m->h1.release(); m->hf.release();
    cudaStreamDestroy(m->stream);
    delete m;
}

void* orbm_stream(orbm_t* m) { return m ? (void*)m->stream : nullptr; }
int orbm_synchronize(orbm_t* m) {
    if (!m) return ORBX_E_INVALID;
    CKM(cudaSetDevice(m->device));
    CKM(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

int orbm_descriptor_distance(const uint8_t* a, const uint8_t* b) {
    uint32_t wa[8], wb[8];
    std::memcpy(wa, a, 32); std::memcpy(wb, b, 32);
    int d = 0;
    for (int i = 0; i < 8; i++) d += __builtin_popcount(wa[i] ^ wb[i]);
    return d;
}

void orbm_three_maxima(const int32_t* counts, int L, int* ind1, int* ind2, int* ind3) {
    int max1 = 0, max2 = 0, max3 = 0;          // ind1..3 keep the caller's initial values (-1) when nothing beats 0
    for (int i = 0; i < L; i++) {
        const int s = counts[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; *ind3 = *ind2; *ind2 = *ind1; *ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; *ind3 = *ind2; *ind2 = i; }
        else if (s > max3) { max3 = s; *ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { *ind2 = -1; *ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { *ind3 = -1; }
}

// caller holds m->mtx
static int bf2nn_device_locked(orbm_t* m, const uint8_t* d_queries, int nq, const uint8_t* d_db, size_t ndb, uint32_t idx_base,
                               orbm_top2* d_out) {
    const int qtiles = (nq + BF_QTILE - 1) / BF_QTILE;
    // enough slices for ~4 resident blocks per SM, at least one 128-row tile each, at most 2^23 rows each
    int slices = std::max(1, (m->numSMs * 4 + qtiles - 1) / qtiles);
    const size_t maxUseful = std::max<size_t>(1, (ndb + BF_ROWS - 1) / BF_ROWS);
    if ((size_t)slices > maxUseful) slices = (int)maxUseful;
    size_t rowsPerSlice = (ndb + slices - 1) / std::max(slices, 1);
    rowsPerSlice = std::max<size_t>(BF_ROWS, (rowsPerSlice + BF_ROWS - 1) / BF_ROWS * BF_ROWS);
    if (rowsPerSlice > (1u << 23)) { rowsPerSlice = 1u << 23; }
    slices = (int)std::max<size_t>(1, (ndb + rowsPerSlice - 1) / rowsPerSlice);
    int rc = m->part.reserve((size_t)slices * nq * sizeof(Top2Key));
    if (rc) return rc;
    bf2nn_kernel<<<dim3(slices, qtiles), BF_THREADS, 0, m->stream>>>((const uint32_t*)d_queries, nq, (const uint32_t*)d_db,
                                                                       (unsigned long long)ndb, (unsigned)rowsPerSlice,
                                                                       idx_base, (Top2Key*)m->part.p);
    merge_slices_kernel<<<(nq + 255) / 256, 256, 0, m->stream>>>((const Top2Key*)m->part.p, slices, nq, d_out);
    CKM(cudaGetLastError());
    return ORBX_OK;
}

int orbm_bf2nn_device(orbm_t* m, const uint8_t* d_queries, int nq, const uint8_t* d_db, size_t ndb, uint32_t idx_base,
                      orbm_top2* d_out) {
    if (!m || !d_queries || nq <= 0 || !d_out || (ndb && !d_db)) return ORBX_E_INVALID;
    if (((uintptr_t)d_queries & 15) || ((uintptr_t)d_db & 15)) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    return bf2nn_device_locked(m, d_queries, nq, d_db, ndb, idx_base, d_out);
}

int orbm_merge_top2_device(orbm_t* m, const orbm_top2* d_parts, int nparts, int nq, orbm_top2* d_out) {
    if (!m || !d_parts || nparts <= 0 || nq <= 0 || !d_out) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    merge_parts_kernel<<<(nq + 255) / 256, 256, 0, m->stream>>>(d_parts, nparts, nq, d_out);
    CKM(cudaGetLastError());
    return ORBX_OK;
}

// One lock over staging, kernels and readback: m->q / m->out / m->part are shared by every call on this context.
int orbm_bf2nn(orbm_t* m, const uint8_t* queries, int nq, const uint8_t* d_db, size_t ndb, uint32_t idx_base, orbm_top2* out) {
    if (!m || !queries || nq <= 0 || !out || (ndb && !d_db)) return ORBX_E_INVALID;
    if ((uintptr_t)d_db & 15) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    int rc = m->q.reserve((size_t)nq * 32); if (rc) return rc;
    rc = m->out.reserve((size_t)nq * sizeof(orbm_top2)); if (rc) return rc;
    CKM(cudaMemcpyAsync(m->q.p, queries, (size_t)nq * 32, cudaMemcpyHostToDevice, m->stream));
    rc = bf2nn_device_locked(m, (const uint8_t*)m->q.p, nq, d_db, ndb, idx_base, (orbm_top2*)m->out.p);
    if (rc) return rc;
    CKM(cudaMemcpyAsync(out, m->out.p, (size_t)nq * sizeof(orbm_top2), cudaMemcpyDeviceToHost, m->stream));
    CKM(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

// ---- orbm_bf2nn_device without any allocation (include/orbm.h: no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops).  The per-slice
// top-2 records bf2nn_kernel hands to merge_slices_kernel live in a file-scope device array instead of the context's
// growable buffer.  Every GPU that loads the library holds the array until exit; every call shares it (calls that
// overlap on one GPU race); nothing in it needs clearing: a call writes every record it later reads.
static __device__ __align__(256) Top2Key g_noalloc_part[ORBM_NOALLOC_PART_RECORDS];

// device address of the array: the name above is a host-side stand-in in host code, not a device pointer
static int noalloc_part_ptr(int device, Top2Key** out) {
    static std::mutex mtx;
    static Top2Key* ptr[64] = {};
    if (device < 0 || device >= 64) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(mtx);
    if (!ptr[device]) {                                  // once per device, with that device current (the caller set it)
        Top2Key* p = nullptr;
        CKM(cudaGetSymbolAddress((void**)&p, g_noalloc_part));
        ptr[device] = p;
    }
    *out = ptr[device];
    return ORBX_OK;
}

int orbm_bf2nn_device_noalloc(orbm_t* m, const uint8_t* d_queries, int nq, const uint8_t* d_db, size_t ndb, uint32_t idx_base,
                              orbm_top2* d_out) {
    if (!m || !d_queries || nq <= 0 || !d_out || (ndb && !d_db)) return ORBX_E_INVALID;
    if (((uintptr_t)d_queries & 15) || ((uintptr_t)d_db & 15)) return ORBX_E_INVALID;
    if (nq > ORBM_NOALLOC_MAX_QUERIES) {
        std::fprintf(stderr, "orbm_bf2nn_device_noalloc: %d queries exceed the fixed scratch (at most %d per call; use orbm_bf2nn_device)\n",
                     nq, ORBM_NOALLOC_MAX_QUERIES);
        return ORBX_E_CAPACITY;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    Top2Key* part = nullptr;
    int rc = noalloc_part_ptr(m->device, &part);
    if (rc) return rc;
    // the slicing of orbm_bf2nn_device, additionally bounded by the scratch: slices * nq records
    const int qtiles = (nq + BF_QTILE - 1) / BF_QTILE;
    int slices = std::max(1, (m->numSMs * 4 + qtiles - 1) / qtiles);
    slices = std::min(slices, ORBM_NOALLOC_PART_RECORDS / nq);
    const size_t maxUseful = std::max<size_t>(1, (ndb + BF_ROWS - 1) / BF_ROWS);
    if ((size_t)slices > maxUseful) slices = (int)maxUseful;
    size_t rowsPerSlice = (ndb + slices - 1) / std::max(slices, 1);
    rowsPerSlice = std::max<size_t>(BF_ROWS, (rowsPerSlice + BF_ROWS - 1) / BF_ROWS * BF_ROWS);
    if (rowsPerSlice > (1u << 23)) rowsPerSlice = 1u << 23;
    slices = (int)std::max<size_t>(1, (ndb + rowsPerSlice - 1) / rowsPerSlice);
    if ((size_t)slices * nq > (size_t)ORBM_NOALLOC_PART_RECORDS) {
        std::fprintf(stderr, "orbm_bf2nn_device_noalloc: %zu rows x %d queries need %d slices, beyond the fixed scratch "
                             "(at most %zu rows at this query count; use orbm_bf2nn_device)\n",
                     ndb, nq, slices, (size_t)(ORBM_NOALLOC_PART_RECORDS / nq) << 23);
        return ORBX_E_CAPACITY;
    }
    bf2nn_kernel<<<dim3(slices, qtiles), BF_THREADS, 0, m->stream>>>((const uint32_t*)d_queries, nq, (const uint32_t*)d_db,
                                                                       (unsigned long long)ndb, (unsigned)rowsPerSlice, idx_base, part);
    merge_slices_kernel<<<(nq + 255) / 256, 256, 0, m->stream>>>(part, slices, nq, d_out);
    CKM(cudaGetLastError());
    return ORBX_OK;
}

int orbm_accept_low_ratio(orbm_top2 r, float nnratio) {
    return r.d1 <= ORBM_TH_LOW && (float)r.d1 < nnratio * (float)r.d2;
}

// thread-instructions per second of one integer instruction, measured on this GPU: which = 0 POPC, 1 LOP3
static int pipe_peak(orbm_t* m, int which, double* per_s) {
    if (!m || !per_s) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    int rc = m->out.reserve(64); if (rc) return rc;
    cudaEvent_t e0, e1;
    CKM(cudaEventCreate(&e0)); CKM(cudaEventCreate(&e1));
    const int iters = 4096, blocks = m->numSMs * 8, threads = 256;
    auto launch = [&](int it, unsigned seed) {
        if (which == 0) popc_peak_kernel<<<blocks, threads, 0, m->stream>>>((unsigned*)m->out.p, it, seed);
        else if (which == 2) mix_peak_kernel<<<blocks, threads, 0, m->stream>>>((unsigned*)m->out.p, it, seed);
        else lop3_peak_kernel<<<blocks, threads, 0, m->stream>>>((unsigned*)m->out.p, it, seed);
    };
    launch(64, 1u);
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        CKM(cudaEventRecord(e0, m->stream));
        launch(iters, 7u + rep);
        CKM(cudaEventRecord(e1, m->stream));
        CKM(cudaEventSynchronize(e1));
        float ms; CKM(cudaEventElapsedTime(&ms, e0, e1));
        best = std::min(best, ms);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *per_s = (double)blocks * threads * iters * (which == 0 ? 8 : which == 2 ? 4 : 16) / (best * 1e-3);
    return ORBX_OK;
}

int orbm_popc_peak(orbm_t* m, double* popc_per_s) { return pipe_peak(m, 0, popc_per_s); }
int orbm_lop3_peak(orbm_t* m, double* lop3_per_s) { return pipe_peak(m, 1, lop3_per_s); }
int orbm_pair_mix_peak(orbm_t* m, double* pairs_per_s) { return pipe_peak(m, 2, pairs_per_s); }

}  // extern "C"

#include "window_host.inl"
#include "bow_host.inl"
#include "vocab_host.inl"
#include "shard_host.inl"
