// Sharded keyframe-descriptor database (BASELINE.json configs[3]): host side, included by matcher.cu.
// One orbm_shard per GPU (one process or thread each).  A query is: bf2nn_kernel over this GPU's rows (global row
// indices through idx_base) -> ncclAllGather of the per-shard (i1, d1, d2) records, nq x 8 bytes per rank, on the
// context stream -> merge_parts_kernel.  The merge is exact: top-2 is decomposable and ties resolve on the global row
// index, i.e. the order of the reference's ascending strict-'<' scan (src/ORBmatcher.cc:205-226).
// NCCL is resolved at run time (dlopen of libnccl.so.2: inside a PyTorch process that is the copy torch already
// loaded, in a plain C++ process the system one), so the library has no link-time NCCL dependency and single-GPU users
// never touch it.  The relocalisation caller this serves is C++ (src/Tracking.cc:1393-1554).

#include <dlfcn.h>
#include <nccl.h>

namespace {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    bool ok = false;
};

const NcclApi& nccl_api() {
    static NcclApi api = [] {
        NcclApi a;
        void* h = nullptr;
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (h) break;
        }
        if (!h) { std::fprintf(stderr, "orbm: cannot load NCCL (%s)\n", dlerror()); return a; }
        a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(h, "ncclGetUniqueId");
        a.CommInitRank = (decltype(a.CommInitRank))dlsym(h, "ncclCommInitRank");
        a.CommDestroy = (decltype(a.CommDestroy))dlsym(h, "ncclCommDestroy");
        a.AllGather = (decltype(a.AllGather))dlsym(h, "ncclAllGather");
        a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
        a.GetVersion = (decltype(a.GetVersion))dlsym(h, "ncclGetVersion");
        a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllGather && a.GetErrorString;
        return a;
    }();
    return api;
}

#define CKN(expr)                                                                                                       \
    do {                                                                                                                \
        ncclResult_t r_ = (expr);                                                                                       \
        if (r_ != ncclSuccess) {                                                                                        \
            std::fprintf(stderr, "orbm: %s failed: %s (%s:%d)\n", #expr, nccl_api().GetErrorString(r_), __FILE__, __LINE__); \
            return ORBX_E_INTERNAL;                                                                                     \
        }                                                                                                               \
    } while (0)

}  // namespace

struct orbm_shard {
    orbm_context* m = nullptr;
    int rank = 0, world = 1;
    ncclComm_t comm = nullptr;
    bool ownComm = false;
    const uint8_t* rows = nullptr;      // device, global rows [idxBase, idxBase + n)
    size_t n = 0;
    uint32_t idxBase = 0;
    DevBuf owned, local, parts, q, out;
    PinBuf hq, hout;
};

extern "C" {

int orbm_nccl_unique_id(uint8_t* id128) {
    if (!id128) return ORBX_E_INVALID;
    const NcclApi& a = nccl_api();
    if (!a.ok) return ORBX_E_UNSUPPORTED;
    ncclUniqueId id;
    CKN(a.GetUniqueId(&id));
    static_assert(sizeof(id) == ORBM_NCCL_ID_BYTES, "ncclUniqueId size");
    std::memcpy(id128, &id, sizeof(id));
    return ORBX_OK;
}

int orbm_nccl_version(int* version) {
    const NcclApi& a = nccl_api();
    if (!version || !a.ok || !a.GetVersion) return ORBX_E_UNSUPPORTED;
    CKN(a.GetVersion(version));
    return ORBX_OK;
}

int orbm_shard_create(orbm_shard_t** out, orbm_t* m, const uint8_t* id128, int rank, int world) {
    if (!out || !m || world < 1 || rank < 0 || rank >= world || (world > 1 && !id128)) return ORBX_E_INVALID;
    if (world > 1 && !nccl_api().ok) return ORBX_E_UNSUPPORTED;
    if (world > 1) CKM(cudaSetDevice(m->device));
    orbm_shard* s = new orbm_shard();
    s->m = m; s->rank = rank; s->world = world;
    if (world > 1) {
        const NcclApi& a = nccl_api();
        ncclUniqueId id;
        std::memcpy(&id, id128, sizeof(id));
        ncclResult_t r = a.CommInitRank(&s->comm, world, id, rank);      // collective over all ranks
        if (r != ncclSuccess) {
            std::fprintf(stderr, "orbm: ncclCommInitRank failed: %s\n", a.GetErrorString(r));
            delete s;
            return ORBX_E_INTERNAL;
        }
        s->ownComm = true;
    }
    *out = s;
    return ORBX_OK;
}

int orbm_shard_create_with_comm(orbm_shard_t** out, orbm_t* m, void* nccl_comm, int rank, int world) {
    if (!out || !m || world < 1 || rank < 0 || rank >= world || (world > 1 && !nccl_comm)) return ORBX_E_INVALID;
    if (world > 1 && !nccl_api().ok) return ORBX_E_UNSUPPORTED;
    orbm_shard* s = new orbm_shard();
    s->m = m; s->rank = rank; s->world = world; s->comm = (ncclComm_t)nccl_comm; s->ownComm = false;
    *out = s;
    return ORBX_OK;
}

void orbm_shard_destroy(orbm_shard_t* s) {
    if (!s) return;
    cudaSetDevice(s->m->device);
    cudaStreamSynchronize(s->m->stream);
    if (s->comm && s->ownComm) nccl_api().CommDestroy(s->comm);
    for (DevBuf* b : {&s->owned, &s->local, &s->parts, &s->q, &s->out}) b->release();
    s->hq.release(); s->hout.release();
    delete s;
}

int orbm_shard_set_rows(orbm_shard_t* s, const uint8_t* d_rows, size_t n, uint32_t idx_base) {
    if (!s || (n && !d_rows) || ((uintptr_t)d_rows & 15)) return ORBX_E_INVALID;
    if ((unsigned long long)idx_base + n > 0xffffffffull) return ORBX_E_INVALID;      // global row indices are 32 bit
    std::lock_guard<std::mutex> lk(s->m->mtx);
    s->rows = d_rows; s->n = n; s->idxBase = idx_base;
    return ORBX_OK;
}

int orbm_shard_upload_rows(orbm_shard_t* s, const uint8_t* rows, size_t n, uint32_t idx_base) {
    if (!s || (n && !rows)) return ORBX_E_INVALID;
    if ((unsigned long long)idx_base + n > 0xffffffffull) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(s->m->mtx);
    CKM(cudaSetDevice(s->m->device));
    CKM(cudaStreamSynchronize(s->m->stream));
    int rc = s->owned.reserve(std::max<size_t>(n, 1) * 32); if (rc) return rc;
    if (n) CKM(cudaMemcpy(s->owned.p, rows, n * 32, cudaMemcpyHostToDevice));
    s->rows = (const uint8_t*)s->owned.p; s->n = n; s->idxBase = idx_base;
    return ORBX_OK;
}

// caller holds the context lock
static int shard_query_locked(orbm_shard* s, const uint8_t* d_queries, int nq, orbm_top2* d_out) {
    orbm_context* m = s->m;
    if (s->world == 1) return bf2nn_device_locked(m, d_queries, nq, s->rows, s->n, s->idxBase, d_out);
    int rc = s->local.reserve((size_t)nq * sizeof(orbm_top2)); if (rc) return rc;
    rc = s->parts.reserve((size_t)s->world * nq * sizeof(orbm_top2)); if (rc) return rc;
    rc = bf2nn_device_locked(m, d_queries, nq, s->rows, s->n, s->idxBase, (orbm_top2*)s->local.p);
    if (rc) return rc;
    CKN(nccl_api().AllGather(s->local.p, s->parts.p, (size_t)nq * sizeof(orbm_top2), ncclUint8, s->comm, m->stream));
    merge_parts_kernel<<<(nq + 255) / 256, 256, 0, m->stream>>>((const orbm_top2*)s->parts.p, s->world, nq, d_out);
    CKM(cudaGetLastError());
    return ORBX_OK;
}

int orbm_shard_bf2nn_device(orbm_shard_t* s, const uint8_t* d_queries, int nq, orbm_top2* d_out) {
    if (!s || !d_queries || nq <= 0 || !d_out || ((uintptr_t)d_queries & 15)) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(s->m->mtx);
    CKM(cudaSetDevice(s->m->device));
    return shard_query_locked(s, d_queries, nq, d_out);
}

int orbm_shard_bf2nn(orbm_shard_t* s, const uint8_t* queries, int nq, orbm_top2* out) {
    if (!s || !queries || nq <= 0 || !out) return ORBX_E_INVALID;
    orbm_context* m = s->m;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    int rc = s->q.reserve((size_t)nq * 32); if (rc) return rc;
    rc = s->out.reserve((size_t)nq * sizeof(orbm_top2)); if (rc) return rc;
    rc = s->hq.reserve((size_t)nq * 32); if (rc) return rc;
    rc = s->hout.reserve((size_t)nq * sizeof(orbm_top2)); if (rc) return rc;
    std::memcpy(s->hq.p, queries, (size_t)nq * 32);                       // pinned staging: the copies stay asynchronous
    CKM(cudaMemcpyAsync(s->q.p, s->hq.p, (size_t)nq * 32, cudaMemcpyHostToDevice, m->stream));
    rc = shard_query_locked(s, (const uint8_t*)s->q.p, nq, (orbm_top2*)s->out.p);
    if (rc) return rc;
    CKM(cudaMemcpyAsync(s->hout.p, s->out.p, (size_t)nq * sizeof(orbm_top2), cudaMemcpyDeviceToHost, m->stream));
    CKM(cudaStreamSynchronize(m->stream));
    std::memcpy(out, s->hout.p, (size_t)nq * sizeof(orbm_top2));
    return ORBX_OK;
}

}  // extern "C"
