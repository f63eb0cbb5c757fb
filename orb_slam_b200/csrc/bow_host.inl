// Host side of ORBmatcher::SearchByBoW (included by matcher.cu).  The device produces every DescriptorDistance the
// reference would evaluate (one row per map point of side 1, over the features of side 2 in the same vocabulary
// node); the scan order, the "already matched" skips, the TH_LOW / mfNNratio acceptance and the rotation histogram
// are order-dependent and are replayed here exactly as written in src/ORBmatcher.cc:159-288 and :590-723.

extern "C" int orbm_search_by_bow(orbm_t* m, int variant,
                                  const uint8_t* desc1, const float* angle1, const uint8_t* valid1, int n1, const orbm_featvec* fv1,
                                  const uint8_t* desc2, const float* angle2, const uint8_t* valid2, int n2, const orbm_featvec* fv2,
                                  float nnratio, int checkOrientation, int32_t* match, int* nmatches) {
    if (!m || !fv1 || !fv2 || !match || !nmatches || n1 < 0 || n2 < 0 || (variant != 0 && variant != 1)) return ORBX_E_INVALID;
    if ((n1 > 0 && (!desc1 || !valid1 || !angle1)) || (n2 > 0 && (!desc2 || !angle2)) || (variant == 1 && n2 > 0 && !valid2))
        return ORBX_E_INVALID;
    if (fv1->nnodes < 0 || fv2->nnodes < 0) return ORBX_E_INVALID;
    {   // well-formed feature vectors: ascending node ids, monotone ranges, indices inside the frames
        const orbm_featvec* fv[2] = {fv1, fv2};
        const int nn[2] = {n1, n2};
        for (int s = 0; s < 2; s++) {
            if (fv[s]->nnodes > 0 && (!fv[s]->node_ids || !fv[s]->start)) return ORBX_E_INVALID;
            if (fv[s]->nnodes == 0) continue;
            if (fv[s]->start[0] != 0) return ORBX_E_INVALID;
            for (int k = 0; k < fv[s]->nnodes; k++) {
                if (k && fv[s]->node_ids[k] <= fv[s]->node_ids[k - 1]) return ORBX_E_INVALID;
                if (fv[s]->start[k + 1] < fv[s]->start[k]) return ORBX_E_INVALID;
            }
            const int total = fv[s]->start[fv[s]->nnodes];
            if (total > 0 && !fv[s]->items) return ORBX_E_INVALID;
            for (int t = 0; t < total; t++) if (fv[s]->items[t] >= (uint32_t)nn[s]) return ORBX_E_INVALID;
        }
    }
    const int nout = variant == ORBM_BOW_KEYFRAME_FRAME ? n2 : n1;
    for (int i = 0; i < nout; i++) match[i] = -1;
    *nmatches = 0;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));

    // ---- the merge walk over the two sorted node lists (:176-249, :614-690); lower_bound on a std::map == first id >= key
    struct Block { int k1, k2; };
    std::vector<Block> blocks;
    {
        int a = 0, b = 0;
        while (a < fv1->nnodes && b < fv2->nnodes) {
            const uint32_t ia = fv1->node_ids[a], ib = fv2->node_ids[b];
            if (ia == ib) { blocks.push_back({a, b}); a++; b++; }
            else if (ia < ib) a = (int)(std::lower_bound(fv1->node_ids + a, fv1->node_ids + fv1->nnodes, ib) - fv1->node_ids);
            else b = (int)(std::lower_bound(fv2->node_ids + b, fv2->node_ids + fv2->nnodes, ia) - fv2->node_ids);
        }
    }
    // ---- one device row per side-1 feature whose map point exists (:192-197, :626-630)
    std::vector<BowRow> rows;
    size_t total = 0;
    for (const Block& bk : blocks) {
        const int s1 = fv1->start[bk.k1], e1 = fv1->start[bk.k1 + 1], s2 = fv2->start[bk.k2], e2 = fv2->start[bk.k2 + 1];
        for (int t = s1; t < e1; t++) {
            const uint32_t i1 = fv1->items[t];
            if (!valid1[i1] || e2 == s2) continue;
            if (total + (size_t)(e2 - s2) > 0xffffffffull) return ORBX_E_CAPACITY;
            rows.push_back(BowRow{i1, s2, e2 - s2, (uint32_t)total});
            total += (size_t)(e2 - s2);
        }
    }
    const uint16_t* dist = nullptr;                      // the distances, in the context's page-locked result buffer
    if (!rows.empty()) {
        // one page-locked block in, one out: [desc1 | desc2 | items2 | rows] -> device, distances back
        const size_t nitems2 = (size_t)fv2->start[fv2->nnodes];      // > 0: some row has candidates
        auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
        const size_t o1 = 0, o2 = o1 + al((size_t)n1 * 32), o3 = o2 + al((size_t)n2 * 32), o4 = o3 + al(nitems2 * 4);
        const size_t inBytes = o4 + al(rows.size() * sizeof(BowRow));
        int rc;
        if ((rc = m->h0.reserve(inBytes))) return rc;
        if ((rc = m->h1.reserve(total * 2))) return rc;
        if ((rc = m->w0.reserve(inBytes))) return rc;
        if ((rc = m->w4.reserve(total * 2))) return rc;
        uint8_t* hb = (uint8_t*)m->h0.p;
        std::memcpy(hb + o1, desc1, (size_t)n1 * 32);
        std::memcpy(hb + o2, desc2, (size_t)n2 * 32);
        std::memcpy(hb + o3, fv2->items, nitems2 * 4);
        std::memcpy(hb + o4, rows.data(), rows.size() * sizeof(BowRow));
        uint8_t* db = (uint8_t*)m->w0.p;
        CKM(cudaMemcpyAsync(db, hb, inBytes, cudaMemcpyHostToDevice, m->stream));
        const int nrows = (int)rows.size();
        bow_dist_kernel<<<(nrows + 7) / 8, 256, 0, m->stream>>>((const BowRow*)(db + o4), nrows, (const uint32_t*)(db + o1),
                                                                 (const uint32_t*)(db + o2), (const uint32_t*)(db + o3), (uint16_t*)m->w4.p);
        CKM(cudaGetLastError());
        CKM(cudaMemcpyAsync(m->h1.p, m->w4.p, total * 2, cudaMemcpyDeviceToHost, m->stream));
        CKM(cudaStreamSynchronize(m->stream));
        dist = (const uint16_t*)m->h1.p;
    }
    // ---- replay
    std::vector<int> rotHist[ORBM_HISTO_LENGTH];
    std::vector<uint8_t> matched2(variant == ORBM_BOW_KEYFRAME_KEYFRAME ? (size_t)n2 : 0, 0);    // vbMatched2, :603
    int nm = 0;
    for (const BowRow& r : rows) {
        int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
        const uint16_t* d = dist + r.out;
        for (int j = 0; j < r.cnt2; j++) {
            const uint32_t i2 = fv2->items[r.off2 + j];
            if (variant == ORBM_BOW_KEYFRAME_FRAME) { if (match[i2] >= 0) continue; }          // :209-210
            else { if (matched2[i2] || !valid2[i2]) continue; }                                 // :646-650
            const int dd = d[j];
            if (dd < bestDist1) { bestDist2 = bestDist1; bestDist1 = dd; bestIdx2 = (int)i2; }
            else if (dd < bestDist2) bestDist2 = dd;
        }
        const bool low = variant == ORBM_BOW_KEYFRAME_FRAME ? bestDist1 <= ORBM_TH_LOW : bestDist1 < ORBM_TH_LOW;   // :228 vs :668
        if (low && (float)bestDist1 < nnratio * (float)bestDist2) {
            if (variant == ORBM_BOW_KEYFRAME_FRAME) match[bestIdx2] = (int32_t)r.idx1;
            else { match[r.idx1] = bestIdx2; matched2[bestIdx2] = 1; }
            if (checkOrientation)
                rotHist[rot_bin(angle1[r.idx1], angle2[bestIdx2])].push_back(variant == ORBM_BOW_KEYFRAME_FRAME ? bestIdx2 : (int)r.idx1);
            nm++;
        }
    }
    if (checkOrientation) {
        int32_t counts[ORBM_HISTO_LENGTH];
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) counts[i] = (int32_t)rotHist[i].size();
        int ind1, ind2, ind3;
        orbm_three_maxima(counts, ORBM_HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { match[idx] = -1; nm--; }
        }
    }
    *nmatches = nm;
    return ORBX_OK;
}

namespace {
// well-formed feature vector: ascending node ids, monotone ranges, indices below n
bool featvec_ok(const orbm_featvec* fv, int n) {
    if (!fv || fv->nnodes < 0) return false;
    if (fv->nnodes == 0) return true;
    if (!fv->node_ids || !fv->start || fv->start[0] != 0) return false;
    for (int k = 0; k < fv->nnodes; k++) {
        if (k && fv->node_ids[k] <= fv->node_ids[k - 1]) return false;
        if (fv->start[k + 1] < fv->start[k]) return false;
    }
    const int total = fv->start[fv->nnodes];
    if (total > 0 && !fv->items) return false;
    for (int t = 0; t < total; t++) if (fv->items[t] >= (uint32_t)n) return false;
    return true;
}
}  // namespace

extern "C" int orbm_search_for_triangulation(orbm_t* m, const orbm_keyframe* kf1, const orbm_keyframe* kf2, const float* Cw1,
                                             const float* R2w, const float* t2w, float fx2, float fy2, float cx2, float cy2,
                                             const float* scaleFactors2, const float* levelSigma2, int nlevels, const float* F12,
                                             int onlyStereo, int checkOri, int32_t* matches12, int* nmatchesOut) {
    if (!m || !kf1 || !kf2 || !Cw1 || !R2w || !t2w || !scaleFactors2 || !levelSigma2 || nlevels <= 0 || !F12 || !matches12 || !nmatchesOut)
        return ORBX_E_INVALID;
    const int n1 = kf1->N, n2 = kf2->N;
    if (n1 < 0 || n2 < 0 || !featvec_ok(kf1->featvec, n1) || !featvec_ok(kf2->featvec, n2)) return ORBX_E_INVALID;
    if ((n1 > 0 && (!kf1->keysUn || !kf1->descriptors || !kf1->uRight || !kf1->has_point)) ||
        (n2 > 0 && (!kf2->keysUn || !kf2->descriptors || !kf2->uRight || !kf2->has_point))) return ORBX_E_INVALID;
    for (int i = 0; i < n2; i++) if (kf2->keysUn[i].octave < 0 || kf2->keysUn[i].octave >= nlevels) return ORBX_E_INVALID;
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    *nmatchesOut = 0;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    const orbm_featvec* fv1 = kf1->featvec; const orbm_featvec* fv2 = kf2->featvec;
    // rows: features of key frame 1 without a map point (and stereo when bOnlyStereo) in nodes both frames share (:758-777)
    std::vector<BowRow> rows;
    {
        int a = 0, b = 0;
        while (a < fv1->nnodes && b < fv2->nnodes) {
            const uint32_t ia = fv1->node_ids[a], ib = fv2->node_ids[b];
            if (ia == ib) {
                const int s2 = fv2->start[b], c2 = fv2->start[b + 1] - s2;
                if (c2 > 0xffff) return ORBX_E_CAPACITY;
                if (c2 > 0)
                    for (int t = fv1->start[a]; t < fv1->start[a + 1]; t++) {
                        const uint32_t i1 = fv1->items[t];
                        if (kf1->has_point[i1]) continue;
                        if (onlyStereo && !(kf1->uRight[i1] >= 0)) continue;
                        rows.push_back(BowRow{i1, s2, c2, 0u});
                    }
                a++; b++;
            } else if (ia < ib) a = (int)(std::lower_bound(fv1->node_ids + a, fv1->node_ids + fv1->nnodes, ib) - fv1->node_ids);
            else b = (int)(std::lower_bound(fv2->node_ids + b, fv2->node_ids + fv2->nnodes, ia) - fv2->node_ids);
        }
    }
    if (rows.empty()) return ORBX_OK;
    // epipole in the second image (:733-740): C2 = R2w*Cw + t2w
    float C2[3];
    mat3_mul_add(R2w, 3, Cw1, t2w, C2);
    const float invz = 1.0f / C2[2];
    TriArgs ta;
    ta.ex = fx2 * C2[0] * invz + cx2;
    ta.ey = fy2 * C2[1] * invz + cy2;
    std::memcpy(ta.F, F12, sizeof(ta.F));
    ta.onlyStereo = onlyStereo ? 1 : 0;
    // device layout in one scratch buffer
    const size_t nitems2 = (size_t)fv2->start[fv2->nnodes];
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    size_t off = 0;
    const size_t oDesc1 = off; off += al((size_t)n1 * 32);
    const size_t oDesc2 = off; off += al((size_t)n2 * 32);
    const size_t oKeys1 = off; off += al((size_t)n1 * sizeof(orbx_keypoint));
    const size_t oKeys2 = off; off += al((size_t)n2 * sizeof(orbx_keypoint));
    const size_t oUr1 = off; off += al((size_t)n1 * 4);
    const size_t oUr2 = off; off += al((size_t)n2 * 4);
    const size_t oHas2 = off; off += al((size_t)n2);
    const size_t oItems2 = off; off += al(nitems2 * 4);
    const size_t oRows = off; off += al(rows.size() * sizeof(BowRow));
    const size_t oSf = off; off += al((size_t)nlevels * 4);
    const size_t oSig = off; off += al((size_t)nlevels * 4);
    const size_t oMatch = off; off += al((size_t)n1 * 4);
    int rc;
    if ((rc = m->w0.reserve(off))) return rc;
    uint8_t* base = (uint8_t*)m->w0.p;
    // every input is staged at its device offset in one page-locked block: one H2D per call
    if ((rc = m->h0.reserve(oMatch))) return rc;
    if ((rc = m->h1.reserve((size_t)std::max(n1, 1) * 4))) return rc;
    uint8_t* hb = (uint8_t*)m->h0.p;
    auto up = [&](size_t o, const void* src, size_t bytes) { if (bytes) std::memcpy(hb + o, src, bytes); };
    up(oDesc1, kf1->descriptors, (size_t)n1 * 32); up(oDesc2, kf2->descriptors, (size_t)n2 * 32);
    up(oKeys1, kf1->keysUn, (size_t)n1 * sizeof(orbx_keypoint)); up(oKeys2, kf2->keysUn, (size_t)n2 * sizeof(orbx_keypoint));
    up(oUr1, kf1->uRight, (size_t)n1 * 4); up(oUr2, kf2->uRight, (size_t)n2 * 4);
    up(oHas2, kf2->has_point, (size_t)n2); up(oItems2, fv2->items, nitems2 * 4);
    up(oRows, rows.data(), rows.size() * sizeof(BowRow));
    up(oSf, scaleFactors2, (size_t)nlevels * 4); up(oSig, levelSigma2, (size_t)nlevels * 4);
    CKM(cudaMemcpyAsync(base, hb, oMatch, cudaMemcpyHostToDevice, m->stream));
    CKM(cudaMemsetAsync(base + oMatch, 0xff, (size_t)n1 * 4, m->stream));                      // -1
    ta.rows = (const BowRow*)(base + oRows); ta.nrows = (int)rows.size();
    ta.desc1 = (const uint32_t*)(base + oDesc1); ta.desc2 = (const uint32_t*)(base + oDesc2); ta.items2 = (const uint32_t*)(base + oItems2);
    ta.keys1 = (const orbx_keypoint*)(base + oKeys1); ta.keys2 = (const orbx_keypoint*)(base + oKeys2);
    ta.uRight1 = (const float*)(base + oUr1); ta.uRight2 = (const float*)(base + oUr2); ta.has2 = base + oHas2;
    ta.scaleFactors2 = (const float*)(base + oSf); ta.levelSigma2 = (const float*)(base + oSig);
    ta.matches12 = (int32_t*)(base + oMatch);
    triangulation_match_kernel<<<(ta.nrows + 7) / 8, 256, 0, m->stream>>>(ta);
    CKM(cudaGetLastError());
    CKM(cudaMemcpyAsync(m->h1.p, base + oMatch, (size_t)n1 * 4, cudaMemcpyDeviceToHost, m->stream));
    CKM(cudaStreamSynchronize(m->stream));
    std::memcpy(matches12, m->h1.p, (size_t)n1 * 4);
    // rotation consistency (:829-875): only the bin sizes matter, every entry of a losing bin is dropped
    int nm = 0;
    std::vector<int> rotHist[ORBM_HISTO_LENGTH];
    for (int i1 = 0; i1 < n1; i1++) {
        if (matches12[i1] < 0) continue;
        nm++;
        if (checkOri) rotHist[rot_bin(kf1->keysUn[i1].angle, kf2->keysUn[matches12[i1]].angle)].push_back(i1);
    }
    if (checkOri) {
        int32_t counts[ORBM_HISTO_LENGTH];
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) counts[i] = (int32_t)rotHist[i].size();
        int ind1 = -1, ind2 = -1, ind3 = -1;
        orbm_three_maxima(counts, ORBM_HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { matches12[idx] = -1; nm--; }
        }
    }
    *nmatchesOut = nm;
    return ORBX_OK;
}
