// Host side of the vocabulary transform (Frame::ComputeBoW) and of the batched
// MapPoint::ComputeDistinctiveDescriptors (included by matcher.cu).  The device does every FORB::distance /
// DescriptorDistance; the std::map bookkeeping of transform() (BowVector::addWeight in feature order, L1/L2
// normalisation in word order, FeatureVector::addFeature) is sequential double arithmetic in the reference and is
// replayed here in the same order (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1126-1194, BowVector.cpp:33-89).

struct orbm_vocab {
    int device = 0;
    int k = 0, L = 0, scoring = 0, weighting = 0, nnodes = 0, nwords = 0;
    std::vector<double> wordWeight;            // Node::weight by word id
    void* dDesc = nullptr; void* dKids = nullptr; void* dNodeId = nullptr; void* dWordId = nullptr;
    VocabDev dev() const { return VocabDev{(const uint4*)dDesc, (const uint2*)dKids, (const uint32_t*)dNodeId, (const uint32_t*)dWordId, L}; }
};

namespace {

int vocab_upload(orbm_vocab* v, const std::vector<uint8_t>& desc, const std::vector<uint2>& kids, const std::vector<uint32_t>& nodeId,
                 const std::vector<uint32_t>& wordId) {
    const size_t n = kids.size();
    CKM(cudaMalloc(&v->dDesc, n * 32)); CKM(cudaMalloc(&v->dKids, n * sizeof(uint2)));
    CKM(cudaMalloc(&v->dNodeId, n * 4)); CKM(cudaMalloc(&v->dWordId, n * 4));
    CKM(cudaMemcpy(v->dDesc, desc.data(), n * 32, cudaMemcpyHostToDevice));
    CKM(cudaMemcpy(v->dKids, kids.data(), n * sizeof(uint2), cudaMemcpyHostToDevice));
    CKM(cudaMemcpy(v->dNodeId, nodeId.data(), n * 4, cudaMemcpyHostToDevice));
    CKM(cudaMemcpy(v->dWordId, wordId.data(), n * 4, cudaMemcpyHostToDevice));
    return ORBX_OK;
}

}  // namespace

extern "C" {

void orbm_vocab_destroy(orbm_vocab_t* v) {
    if (!v) return;
    cudaSetDevice(v->device);
    for (void* p : {v->dDesc, v->dKids, v->dNodeId, v->dWordId}) if (p) cudaFree(p);
    delete v;
}

int orbm_vocab_create(orbm_t* m, int k, int L, int scoring, int weighting, int nnodes, const int32_t* parent,
                      const uint8_t* is_leaf, const uint8_t* desc, const double* weight, orbm_vocab_t** out) {
    if (!m || !out || nnodes < 1 || L < 1 || scoring < 0 || scoring > 5 || weighting < 0 || weighting > 3) return ORBX_E_INVALID;
    if (nnodes > 1 && (!parent || !is_leaf || !desc || !weight)) return ORBX_E_INVALID;
    // children in append order (= ascending node id, TemplatedVocabulary.h:1390-1393), as a CSR
    std::vector<int32_t> cnt((size_t)nnodes + 1, 0);
    for (int i = 1; i < nnodes; i++) {
        if (parent[i] < 0 || parent[i] >= i) return ORBX_E_INVALID;       // a parent line precedes its children's lines
        cnt[(size_t)parent[i] + 1]++;
    }
    for (int i = 0; i < nnodes; i++) {
        if (cnt[(size_t)i + 1] > 255) return ORBX_E_INVALID;
        if (i > 0 && cnt[(size_t)i + 1] == 0 && !is_leaf[i]) return ORBX_E_INVALID;   // childless node that is not a word
        cnt[(size_t)i + 1] += cnt[i];
    }
    std::vector<int32_t> child((size_t)std::max(nnodes - 1, 0)), fill(cnt.begin(), cnt.end() - 1);
    for (int i = 1; i < nnodes; i++) child[(size_t)fill[parent[i]]++] = i;
    // word ids in line order of the leaf lines (:1410-1417)
    std::vector<uint32_t> wordOf((size_t)nnodes, 0);
    orbm_vocab* v = new orbm_vocab();
    v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting; v->nnodes = nnodes; v->device = m->device;
    for (int i = 1; i < nnodes; i++)
        if (is_leaf[i]) { wordOf[i] = (uint32_t)v->wordWeight.size(); v->wordWeight.push_back(weight[i]); }
    v->nwords = (int)v->wordWeight.size();
    // breadth-first slots: the children of a node are consecutive
    std::vector<int32_t> order; order.reserve(nnodes); order.push_back(0);
    std::vector<uint2> kids((size_t)nnodes);
    for (size_t s = 0; s < order.size(); s++) {
        const int node = order[s];
        const int c0 = cnt[node], c1 = cnt[(size_t)node + 1];
        kids[s] = make_uint2((uint32_t)order.size(), (uint32_t)(c1 - c0));
        for (int c = c0; c < c1; c++) order.push_back(child[c]);
    }
    std::vector<uint8_t> sdesc((size_t)nnodes * 32, 0);
    std::vector<uint32_t> nodeId((size_t)nnodes), wordId((size_t)nnodes);
    for (size_t s = 0; s < order.size(); s++) {
        const int node = order[s];
        if (node > 0) std::memcpy(&sdesc[s * 32], desc + (size_t)node * 32, 32);
        nodeId[s] = (uint32_t)node; wordId[s] = wordOf[node];
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    cudaError_t e = cudaSetDevice(m->device);
    int rc = e != cudaSuccess ? (int)e : vocab_upload(v, sdesc, kids, nodeId, wordId);
    if (rc) { orbm_vocab_destroy(v); return rc; }
    *out = v;
    return ORBX_OK;
}

// TemplatedVocabulary::loadFromTextFile (:1338-1426).  Canonical choice: blank lines are skipped - the reference's
// `while(!f.eof())` loop turns the empty read after a final newline into one extra child of the root whose
// descriptor is uninitialised memory, which no implementation can reproduce.
int orbm_vocab_load_text(orbm_t* m, const char* path, orbm_vocab_t** out) {
    if (!m || !path || !out) return ORBX_E_INVALID;
    FILE* f = std::fopen(path, "rb");
    if (!f) return ORBX_E_INVALID;
    std::vector<char> buf;
    {
        char tmp[1 << 16]; size_t got;
        while ((got = std::fread(tmp, 1, sizeof tmp, f)) > 0) buf.insert(buf.end(), tmp, tmp + got);
        std::fclose(f);
        buf.push_back('\n'); buf.push_back(0);
    }
    char* p = buf.data();
    char* end;
    const long k = std::strtol(p, &end, 10); p = end;
    const long L = std::strtol(p, &end, 10); p = end;
    const long n1 = std::strtol(p, &end, 10); p = end;
    const long n2 = std::strtol(p, &end, 10); p = end;
    if (k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) return ORBX_E_INVALID;   // :1359-1363
    while (*p && *p != '\n') p++;
    std::vector<int32_t> parent(1, 0);
    std::vector<uint8_t> leaf(1, 0), desc(32, 0);
    std::vector<double> weight(1, 0.0);
    while (*p) {
        while (*p == '\n' || *p == '\r' || *p == ' ' || *p == '\t') p++;
        if (!*p) break;
        char* eol = p; while (*eol && *eol != '\n') eol++;
        const char save = *eol; *eol = 0;
        const long pid = std::strtol(p, &end, 10); if (end == p) return ORBX_E_INVALID; p = end;
        const long isLeaf = std::strtol(p, &end, 10); if (end == p) return ORBX_E_INVALID; p = end;
        uint8_t d[32];
        for (int i = 0; i < 32; i++) { const long b = std::strtol(p, &end, 10); if (end == p) return ORBX_E_INVALID; p = end; d[i] = (uint8_t)b; }
        const double w = std::strtod(p, &end); if (end == p) return ORBX_E_INVALID;
        parent.push_back((int32_t)pid); leaf.push_back(isLeaf > 0); desc.insert(desc.end(), d, d + 32); weight.push_back(w);
        *eol = save; p = eol;
    }
    return orbm_vocab_create(m, (int)k, (int)L, (int)n1, (int)n2, (int)parent.size(), parent.data(), leaf.data(), desc.data(),
                             weight.data(), out);
}

int orbm_vocab_info(const orbm_vocab_t* v, int* k, int* L, int* nnodes, int* nwords, int* scoring, int* weighting) {
    if (!v) return ORBX_E_INVALID;
    if (k) *k = v->k;
    if (L) *L = v->L;
    if (nnodes) *nnodes = v->nnodes;
    if (nwords) *nwords = v->nwords;
    if (scoring) *scoring = v->scoring;
    if (weighting) *weighting = v->weighting;
    return ORBX_OK;
}

int orbm_vocab_descend_device(orbm_t* m, const orbm_vocab_t* v, const uint8_t* d_desc, size_t n, int levelsup, uint32_t* d_word_node) {
    if (!m || !v || v->device != m->device || (n && (!d_desc || !d_word_node)) || ((uintptr_t)d_desc & 15) || ((uintptr_t)d_word_node & 7))
        return ORBX_E_INVALID;
    if (n == 0) return ORBX_OK;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    const size_t threads = n * VOC_GROUP;
    vocab_descend_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, m->stream>>>(v->dev(), (const uint4*)d_desc, n, levelsup, (uint2*)d_word_node);
    CKM(cudaGetLastError());
    return ORBX_OK;
}

}  // extern "C"

namespace {
// descent of n host descriptors; result in the pinned buffer h1 as {word, node} pairs
int vocab_descend_host(orbm_t* m, const orbm_vocab_t* v, const uint8_t* desc, int n, int levelsup, const uint2** res) {
    int rc;
    if ((rc = m->w0.reserve((size_t)n * 32))) return rc;
    if ((rc = m->w1.reserve((size_t)n * 8))) return rc;
    if ((rc = m->h1.reserve((size_t)n * 8))) return rc;
    if ((rc = m->h0.reserve((size_t)n * 32))) return rc;
    std::memcpy(m->h0.p, desc, (size_t)n * 32);          // page-locked staging: the H2D is a plain DMA
    CKM(cudaMemcpyAsync(m->w0.p, m->h0.p, (size_t)n * 32, cudaMemcpyHostToDevice, m->stream));
    vocab_descend_kernel<<<(unsigned)(((size_t)n * VOC_GROUP + 255) / 256), 256, 0, m->stream>>>(v->dev(), (const uint4*)m->w0.p, (size_t)n,
                                                                                                   levelsup, (uint2*)m->w1.p);
    CKM(cudaGetLastError());
    CKM(cudaMemcpyAsync(m->h1.p, m->w1.p, (size_t)n * 8, cudaMemcpyDeviceToHost, m->stream));
    CKM(cudaStreamSynchronize(m->stream));
    *res = (const uint2*)m->h1.p;
    return ORBX_OK;
}
}  // namespace

extern "C" {

int orbm_vocab_descend(orbm_t* m, const orbm_vocab_t* v, const uint8_t* desc, int n, int levelsup, uint32_t* word_id,
                       uint32_t* node_id, double* weight) {
    if (!m || !v || v->device != m->device || n < 0 || (n && (!desc || !word_id || !node_id))) return ORBX_E_INVALID;
    if (n == 0) return ORBX_OK;
    if (v->nwords == 0) return ORBX_E_INVALID;              // empty(): transform() returns before any descent
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    const uint2* r;
    int rc = vocab_descend_host(m, v, desc, n, levelsup, &r);
    if (rc) return rc;
    for (int i = 0; i < n; i++) {
        word_id[i] = r[i].x; node_id[i] = r[i].y;
        if (weight) weight[i] = v->wordWeight[r[i].x];
    }
    return ORBX_OK;
}

int orbm_vocab_transform(orbm_t* m, const orbm_vocab_t* v, const uint8_t* desc, int n, int levelsup, uint32_t* word_ids,
                         double* word_vals, int* nwords, uint32_t* node_ids, int32_t* node_start, uint32_t* items, int* nnodes) {
    if (!m || !v || v->device != m->device || n < 0 || !nwords || !nnodes || !node_start) return ORBX_E_INVALID;
    if (n && (!desc || !word_ids || !word_vals || !node_ids || !items)) return ORBX_E_INVALID;
    *nwords = 0; *nnodes = 0; node_start[0] = 0;
    if (n == 0 || v->nwords == 0) return ORBX_OK;           // v.clear(); fv.clear(); if(empty()) return; (:1131-1137)
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    const uint2* r;
    int rc = vocab_descend_host(m, v, desc, n, levelsup, &r);
    if (rc) return rc;
    // features that are not stopped (w > 0, :1157/:1185), keyed for the two ordered maps; a stable order inside a key
    // is the feature order, which is what addWeight / addFeature see
    std::vector<uint64_t> wk, nk;
    wk.reserve(n); nk.reserve(n);
    for (int i = 0; i < n; i++) {
        if (!(v->wordWeight[r[i].x] > 0)) continue;
        wk.push_back(((uint64_t)r[i].x << 32) | (uint32_t)i);
        nk.push_back(((uint64_t)r[i].y << 32) | (uint32_t)i);
    }
    std::sort(wk.begin(), wk.end());
    std::sort(nk.begin(), nk.end());
    const bool tf = v->weighting == ORBM_W_TF || v->weighting == ORBM_W_TF_IDF;
    int nw = 0;
    for (size_t a = 0; a < wk.size();) {
        const uint32_t id = (uint32_t)(wk[a] >> 32);
        size_t b = a; while (b < wk.size() && (uint32_t)(wk[b] >> 32) == id) b++;
        const double w = v->wordWeight[id];
        double val = w;                                     // insert(id, w), then "+= w" per further feature (BowVector.cpp:33-45)
        if (tf) for (size_t t = a + 1; t < b; t++) val += w;        // addIfNotExist keeps the first value (:49-57)
        word_ids[nw] = id; word_vals[nw] = val; nw++;
        a = b;
    }
    const bool must = v->scoring != ORBM_S_DOT_PRODUCT;    // ScoringObject.h:74-89
    if (tf && nw > 0 && !must) {                            // :1165-1171
        const double nd = (double)nw;
        for (int t = 0; t < nw; t++) word_vals[t] /= nd;
    }
    if (must) {                                             // BowVector::normalize (BowVector.cpp:61-83)
        double norm = 0.0;
        if (v->scoring != ORBM_S_L2_NORM) { for (int t = 0; t < nw; t++) norm += std::fabs(word_vals[t]); }
        else { for (int t = 0; t < nw; t++) norm += word_vals[t] * word_vals[t]; norm = std::sqrt(norm); }
        if (norm > 0.0) for (int t = 0; t < nw; t++) word_vals[t] /= norm;
    }
    *nwords = nw;
    int nn = 0, pos = 0;
    for (size_t a = 0; a < nk.size();) {
        const uint32_t id = (uint32_t)(nk[a] >> 32);
        node_ids[nn] = id; node_start[nn] = pos; nn++;
        while (a < nk.size() && (uint32_t)(nk[a] >> 32) == id) items[pos++] = (uint32_t)nk[a++];
    }
    node_start[nn] = pos;
    *nnodes = nn;
    return ORBX_OK;
}

int orbm_distinctive_descriptors(orbm_t* m, const uint8_t* desc, const int32_t* start, int npoints, int32_t* best) {
    if (!m || npoints < 0 || (npoints && (!start || !best))) return ORBX_E_INVALID;
    if (npoints == 0) return ORBX_OK;
    if (start[0] != 0) return ORBX_E_INVALID;
    for (int p = 0; p < npoints; p++) if (start[p + 1] < start[p]) return ORBX_E_INVALID;
    const size_t total = (size_t)start[npoints];
    if (total && !desc) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    int rc;
    if ((rc = m->w0.reserve(std::max<size_t>(total, 1) * 32))) return rc;
    if ((rc = m->w1.reserve(((size_t)npoints + 1) * 4))) return rc;
    if ((rc = m->w2.reserve((size_t)npoints * 4))) return rc;
    if (total) CKM(cudaMemcpyAsync(m->w0.p, desc, total * 32, cudaMemcpyHostToDevice, m->stream));
    CKM(cudaMemcpyAsync(m->w1.p, start, ((size_t)npoints + 1) * 4, cudaMemcpyHostToDevice, m->stream));
    distinctive_kernel<<<(npoints + 7) / 8, 256, 0, m->stream>>>((const uint4*)m->w0.p, (const int32_t*)m->w1.p, npoints, (int32_t*)m->w2.p);
    CKM(cudaGetLastError());
    CKM(cudaMemcpyAsync(best, m->w2.p, (size_t)npoints * 4, cudaMemcpyDeviceToHost, m->stream));
    CKM(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

}  // extern "C"
