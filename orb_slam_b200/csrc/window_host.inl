// Host side of the grid / windowed matchers (included by matcher.cu).
// The device produces, per query, the GetFeaturesInArea candidate list with Hamming
// distances; the sequential, order-dependent acceptance rules of the reference are
// replayed here over those lists so results equal the reference exactly.

namespace {

struct DeviceFrame {            // target frame resident on the device (context scratch buffers w0..w3)
    const orbx_keypoint* keysUn; const uint32_t* desc; const int32_t* cellStart; const int32_t* items;
    GridParams g; int N;
};

int upload_frame(orbm_context* m, const orbm_frame* f, DeviceFrame& d) {
    const size_t N = (size_t)std::max(f->N, 1);
    int rc;
    if ((rc = m->w0.reserve(N * sizeof(orbx_keypoint)))) return rc;
    if ((rc = m->w1.reserve(N * 32))) return rc;
    if ((rc = m->w2.reserve((GRID_CELLS + 1) * 4 + N * 4))) return rc;     // cellStart | cellOf
    if ((rc = m->w3.reserve(N * 4))) return rc;                            // items
    if (f->N > 0) {                                    // through the pinned block: pageable sources stall the stream
        const size_t kb = (size_t)f->N * sizeof(orbx_keypoint), db = (size_t)f->N * 32, ko = (kb + 255) & ~(size_t)255;
        if ((rc = m->hf.reserve(ko + db))) return rc;
        CKM(cudaStreamSynchronize(m->stream));         // (the block may still feed the previous call's copies)
        std::memcpy(m->hf.p, f->keysUn, kb); std::memcpy((uint8_t*)m->hf.p + ko, f->descriptors, db);
        CKM(cudaMemcpyAsync(m->w0.p, m->hf.p, kb, cudaMemcpyHostToDevice, m->stream));
        CKM(cudaMemcpyAsync(m->w1.p, (uint8_t*)m->hf.p + ko, db, cudaMemcpyHostToDevice, m->stream));
    }
    d.keysUn = (const orbx_keypoint*)m->w0.p; d.desc = (const uint32_t*)m->w1.p;
    int32_t* cellStart = (int32_t*)m->w2.p; int32_t* cellOf = cellStart + GRID_CELLS + 1;
    d.cellStart = cellStart; d.items = (const int32_t*)m->w3.p; d.N = f->N;
    d.g = GridParams{f->mnMinX, f->mnMinY, f->gridElementWidthInv, f->gridElementHeightInv};
    grid_build_kernel<<<1, 1024, 0, m->stream>>>(d.keysUn, f->N, d.g, cellOf, cellStart, (int32_t*)m->w3.p);
    CKM(cudaGetLastError());
    return 0;
}

// runs the window queries against the uploaded frame; lists[q] = offs[q]..offs[q+1] in `pairs`.
// Count, scan and fill are enqueued back to back against a pair buffer sized from the query count, and offsets and pairs
// come back in one pinned block: ONE host synchronisation per call (a second round only if the buffer was too small).
int run_queries(orbm_context* m, const DeviceFrame& d, const std::vector<WQuery>& queries, std::vector<int32_t>& offs,
                std::vector<int2>& pairs) {
    const int nq = (int)queries.size();
    offs.assign(nq + 1, 0);
    pairs.clear();
    if (nq == 0) return 0;
    int rc;
    if ((rc = m->w4.reserve((size_t)nq * sizeof(WQuery) + (size_t)(2 * nq + 2) * 4))) return rc;
    if ((rc = m->h0.reserve((size_t)nq * sizeof(WQuery)))) return rc;
    WQuery* dq = (WQuery*)m->w4.p;
    int32_t* dcnt = (int32_t*)(dq + nq); int32_t* doff = dcnt + nq;
    std::memcpy(m->h0.p, queries.data(), (size_t)nq * sizeof(WQuery));
    CKM(cudaMemcpyAsync(dq, m->h0.p, (size_t)nq * sizeof(WQuery), cudaMemcpyHostToDevice, m->stream));
    const dim3 grid((nq + 7) / 8);
    window_kernel<false><<<grid, 256, 0, m->stream>>>(dq, nq, d.g, d.keysUn, d.desc, d.cellStart, d.items, dcnt, nullptr, nullptr, 0);
    scan_counts_kernel<<<1, 1024, 0, m->stream>>>(dcnt, nq, doff);
    // sized from the previous call's total (tracking repeats the same searches frame after frame)
    size_t cap = std::max((size_t)nq * 8 + 1024, m->lastPairs + m->lastPairs / 4 + 256);
    for (int round = 0; round < 2; round++) {
        const size_t offBytes = (((size_t)nq + 1) * 4 + 15) & ~(size_t)15;
        if ((rc = m->w5.reserve(cap * sizeof(int2)))) return rc;
        if ((rc = m->h1.reserve(offBytes + cap * sizeof(int2)))) return rc;
        window_kernel<true><<<grid, 256, 0, m->stream>>>(dq, nq, d.g, d.keysUn, d.desc, d.cellStart, d.items, nullptr, doff, (int2*)m->w5.p, (int)cap);
        uint8_t* h = (uint8_t*)m->h1.p;
        CKM(cudaMemcpyAsync(h, doff, (size_t)(nq + 1) * 4, cudaMemcpyDeviceToHost, m->stream));
        CKM(cudaMemcpyAsync(h + offBytes, m->w5.p, cap * sizeof(int2), cudaMemcpyDeviceToHost, m->stream));
        CKM(cudaStreamSynchronize(m->stream));
        std::memcpy(offs.data(), h, (size_t)(nq + 1) * 4);
        const size_t total = (size_t)offs[nq];
        m->lastPairs = total;
        if (total <= cap) {
            pairs.assign((const int2*)(h + offBytes), (const int2*)(h + offBytes) + total);
            break;
        }
        cap = total;                                   // rare: dense windows; second round with the exact size
    }
    CKM(cudaGetLastError());
    return 0;
}

inline void set_desc(WQuery& q, const uint8_t* d) { std::memcpy(q.desc, d, 32); }

// rotation-histogram bin of :528-540 / :1501-1508
inline int rot_bin(float a1, float a2) {
    const float factor = 1.0f / ORBM_HISTO_LENGTH;
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)roundf(rot * factor);
    if (bin == ORBM_HISTO_LENGTH) bin = 0;
    return bin;
}

}  // namespace

extern "C" {

int orbm_grid_build(orbm_t* m, const orbm_frame* f, int32_t* cell_start, int32_t* items) {
    if (!m || !f || !cell_start || !items || f->N < 0) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    DeviceFrame d;
    int rc = upload_frame(m, f, d);
    if (rc) return rc;
    CKM(cudaMemcpyAsync(cell_start, d.cellStart, (GRID_CELLS + 1) * 4, cudaMemcpyDeviceToHost, m->stream));
    if (f->N > 0) CKM(cudaMemcpyAsync(items, d.items, (size_t)f->N * 4, cudaMemcpyDeviceToHost, m->stream));
    CKM(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

int orbm_features_in_area(orbm_t* m, const orbm_frame* f, float x, float y, float r, int minLevel, int maxLevel,
                          int32_t* out, int capacity, int* n_out) {
    if (!m || !f || !n_out) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    DeviceFrame d;
    int rc = upload_frame(m, f, d);
    if (rc) return rc;
    std::vector<WQuery> q(1);
    std::memset(&q[0], 0, sizeof(WQuery));
    q[0].x = x; q[0].y = y; q[0].r = r; q[0].minLevel = minLevel; q[0].maxLevel = maxLevel;
    std::vector<int32_t> offs; std::vector<int2> pairs;
    if ((rc = run_queries(m, d, q, offs, pairs))) return rc;
    *n_out = (int)pairs.size();
    for (int i = 0; i < (int)pairs.size() && i < capacity; i++) out[i] = pairs[i].x;
    return ORBX_OK;
}

// ORBmatcher::SearchForInitialization, src/ORBmatcher.cc:417-588
int orbm_search_for_initialization(orbm_t* m, const orbm_frame* F1, const orbm_frame* F2, float* prev, int32_t* matches12,
                                   int windowSize, float nnratio, int checkOri, int* nmatchesOut) {
    if (!m || !F1 || !F2 || !prev || !matches12 || !nmatchesOut) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    DeviceFrame d;
    int rc = upload_frame(m, F2, d);
    if (rc) return rc;
    std::vector<WQuery> queries; std::vector<int> qOf;
    for (int i1 = 0; i1 < F1->N; i1++) {
        const int level1 = F1->keysUn[i1].octave;
        if (level1 > 0) continue;                                   // :451
        WQuery q; std::memset(&q, 0, sizeof(q));
        q.x = prev[2 * i1]; q.y = prev[2 * i1 + 1]; q.r = (float)windowSize; q.minLevel = level1; q.maxLevel = level1;
        set_desc(q, F1->descriptors + 32 * (size_t)i1);
        queries.push_back(q); qOf.push_back(i1);
    }
    std::vector<int32_t> offs; std::vector<int2> pairs;
    if ((rc = run_queries(m, d, queries, offs, pairs))) return rc;

    int nmatches = 0;
    for (int i = 0; i < F1->N; i++) matches12[i] = -1;
    std::vector<int> rotHist[ORBM_HISTO_LENGTH];
    std::vector<int> matchedDistance(F2->N, INT32_MAX), matches21(F2->N, -1);
    for (size_t qi = 0; qi < queries.size(); qi++) {
        const int i1 = qOf[qi];
        if (offs[qi + 1] == offs[qi]) continue;
        int bestDist = INT32_MAX, bestDist2 = INT32_MAX, bestIdx2 = -1;
        for (int k = offs[qi]; k < offs[qi + 1]; k++) {
            const int i2 = pairs[k].x, dist = pairs[k].y;
            if (matchedDistance[i2] <= dist) continue;             // :488
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= ORBM_TH_LOW) {
            if (bestDist < (float)bestDist2 * nnratio) {
                if (matches21[bestIdx2] >= 0) { matches12[matches21[bestIdx2]] = -1; nmatches--; }
                matches12[i1] = bestIdx2; matches21[bestIdx2] = i1; matchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (checkOri) rotHist[rot_bin(F1->keysUn[i1].angle, F2->keysUn[bestIdx2].angle)].push_back(i1);
            }
        }
    }
    if (checkOri) {
        int32_t counts[ORBM_HISTO_LENGTH];
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) counts[i] = (int)rotHist[i].size();
        int ind1 = -1, ind2 = -1, ind3 = -1;
        orbm_three_maxima(counts, ORBM_HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; nmatches--; }
        }
    }
    for (int i1 = 0; i1 < F1->N; i1++)
        if (matches12[i1] >= 0) { prev[2 * i1] = F2->keysUn[matches12[i1]].x; prev[2 * i1 + 1] = F2->keysUn[matches12[i1]].y; }
    *nmatchesOut = nmatches;
    return ORBX_OK;
}

}  // extern "C"

namespace {
// cv::Mat float algebra as OpenCV 4.13 evaluates it for 3x3 * 3x1 (+ 3x1): products and sums in float, left to
// right, no FMA; the "+ c" of the fused gemm is applied in double (pinned against cv2.gemm in tests/golden).
inline void mat3_mul_add(const float* T, int rowStride, const float* x, const float* c, float* out) {
    for (int i = 0; i < 3; i++) {
        const float* a = T + i * rowStride;
        float t = a[0] * x[0];
        t = t + a[1] * x[1];
        t = t + a[2] * x[2];
        out[i] = c ? (float)((double)t + (double)c[i]) : t;
    }
}
}  // namespace

extern "C" {

// ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono), src/ORBmatcher.cc:1396-1538
int orbm_search_by_projection_frame(orbm_t* m, const orbm_frame* cur, const orbm_frame* last, const orbm_mappoint* lp,
                                    const int32_t* curObs, float th, int bMono, int checkOri, int32_t* curAssign,
                                    int* nmatchesOut) {
    if (!m || !cur || !last || !lp || !curObs || !curAssign || !nmatchesOut) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    DeviceFrame d;
    int rc = upload_frame(m, cur, d);
    if (rc) return rc;
    const float* Tc = cur->Tcw; const float* Tl = last->Tcw;
    const float tcw[3] = {Tc[3], Tc[7], Tc[11]}, tlw[3] = {Tl[3], Tl[7], Tl[11]};
    float twc[3];                                          // -Rcw.t()*tcw (general gemm path: double accumulation)
    for (int i = 0; i < 3; i++) {
        double s = 0;
        for (int k = 0; k < 3; k++) s += (double)Tc[4 * k + i] * (double)tcw[k];
        twc[i] = (float)(s * -1.0);
    }
    float tlc[3];
    mat3_mul_add(Tl, 4, twc, tlw, tlc);
    const bool bForward = tlc[2] > cur->b && !bMono;
    const bool bBackward = -tlc[2] > cur->b && !bMono;

    struct Q { int i; float u, v, radius, invzc; };
    std::vector<Q> meta; std::vector<WQuery> queries;
    for (int i = 0; i < last->N; i++) {
        if (!lp[i].valid || lp[i].outlier) continue;
        float x3Dc[3];
        mat3_mul_add(Tc, 4, lp[i].world, tcw, x3Dc);
        const float xc = x3Dc[0], yc = x3Dc[1];
        const float invzc = (float)(1.0 / x3Dc[2]);
        if (invzc < 0) continue;
        const float u = cur->fx * xc * invzc + cur->cx;
        const float v = cur->fy * yc * invzc + cur->cy;
        if (u < cur->mnMinX || u > cur->mnMaxX) continue;
        if (v < cur->mnMinY || v > cur->mnMaxY) continue;
        const int nLastOctave = last->keys[i].octave;
        const float radius = th * cur->scaleFactors[nLastOctave];
        WQuery q; std::memset(&q, 0, sizeof(q));
        q.x = u; q.y = v; q.r = radius;
        if (bForward) { q.minLevel = nLastOctave; q.maxLevel = -1; }                 // GetFeaturesInArea(u,v,radius,nLastOctave)
        else if (bBackward) { q.minLevel = 0; q.maxLevel = nLastOctave; }
        else { q.minLevel = nLastOctave - 1; q.maxLevel = nLastOctave + 1; }
        set_desc(q, lp[i].descriptor);
        queries.push_back(q); meta.push_back(Q{i, u, v, radius, invzc});
    }
    std::vector<int32_t> offs; std::vector<int2> pairs;
    if ((rc = run_queries(m, d, queries, offs, pairs))) return rc;

    int nmatches = 0;
    std::vector<int> obs(curObs, curObs + cur->N);       // Observations() of the point currently stored, -1 = NULL
    for (int i = 0; i < cur->N; i++) curAssign[i] = -2;
    std::vector<int> rotHist[ORBM_HISTO_LENGTH];
    for (size_t qi = 0; qi < queries.size(); qi++) {
        if (offs[qi + 1] == offs[qi]) continue;
        const Q& mq = meta[qi];
        int bestDist = 256, bestIdx2 = -1;
        for (int k = offs[qi]; k < offs[qi + 1]; k++) {
            const int i2 = pairs[k].x;
            if (obs[i2] > 0) continue;                                        // :1471-1473
            if (cur->uRight && cur->uRight[i2] > 0) {
                const float ur = mq.u - cur->bf * mq.invzc;
                const float er = fabsf(ur - cur->uRight[i2]);
                if (er > mq.radius) continue;
            }
            const int dist = pairs[k].y;
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= ORBM_TH_HIGH) {
            curAssign[bestIdx2] = mq.i; obs[bestIdx2] = lp[mq.i].observations;
            nmatches++;
            if (checkOri) rotHist[rot_bin(last->keysUn[mq.i].angle, cur->keysUn[bestIdx2].angle)].push_back(bestIdx2);
        }
    }
    if (checkOri) {
        int32_t counts[ORBM_HISTO_LENGTH];
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) counts[i] = (int)rotHist[i].size();
        int ind1 = -1, ind2 = -1, ind3 = -1;
        orbm_three_maxima(counts, ORBM_HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) { curAssign[idx] = -1; nmatches--; }
    }
    *nmatchesOut = nmatches;
    return ORBX_OK;
}

// ORBmatcher::SearchByProjection(F, vpMapPoints, th), src/ORBmatcher.cc:45-129
int orbm_search_by_projection_points(orbm_t* m, const orbm_frame* F, const orbm_mappoint* pts, int npts,
                                     const int32_t* frameObs, float th, float nnratio, int32_t* frameAssign,
                                     int* nmatchesOut) {
    if (!m || !F || (!pts && npts) || !frameObs || !frameAssign || !nmatchesOut) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    DeviceFrame d;
    int rc = upload_frame(m, F, d);
    if (rc) return rc;
    const bool bFactor = th != 1.0;
    struct Q { int i; float rs; };
    std::vector<Q> meta; std::vector<WQuery> queries;
    for (int i = 0; i < npts; i++) {
        const orbm_mappoint& p = pts[i];
        if (!p.trackInView) continue;
        if (p.bad) continue;
        const int lvl = p.trackScaleLevel;
        float r = p.trackViewCos > 0.998 ? 2.5f : 4.0f;                        // RadiusByViewingCos, :131-137
        if (bFactor) r *= th;
        WQuery q; std::memset(&q, 0, sizeof(q));
        q.x = p.trackProjX; q.y = p.trackProjY; q.r = r * F->scaleFactors[lvl]; q.minLevel = lvl - 1; q.maxLevel = lvl;
        set_desc(q, p.descriptor);
        queries.push_back(q); meta.push_back(Q{i, r * F->scaleFactors[lvl]});
    }
    std::vector<int32_t> offs; std::vector<int2> pairs;
    if ((rc = run_queries(m, d, queries, offs, pairs))) return rc;
    int nmatches = 0;
    std::vector<int> obs(frameObs, frameObs + F->N);
    for (int i = 0; i < F->N; i++) frameAssign[i] = -2;
    for (size_t qi = 0; qi < queries.size(); qi++) {
        if (offs[qi + 1] == offs[qi]) continue;
        const orbm_mappoint& p = pts[meta[qi].i];
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int k = offs[qi]; k < offs[qi + 1]; k++) {
            const int idx = pairs[k].x;
            if (obs[idx] > 0) continue;                                        // :87-89
            if (F->uRight && F->uRight[idx] > 0) {
                const float er = fabsf(p.trackProjXR - F->uRight[idx]);
                if (er > meta[qi].rs) continue;
            }
            const int dist = pairs[k].y;
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = F->keysUn[idx].octave; bestIdx = idx;
            } else if (dist < bestDist2) { bestLevel2 = F->keysUn[idx].octave; bestDist2 = dist; }
        }
        if (bestDist <= ORBM_TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
            frameAssign[bestIdx] = meta[qi].i; obs[bestIdx] = p.observations;
            nmatches++;
        }
    }
    *nmatchesOut = nmatches;
    return ORBX_OK;
}

// ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist), src/ORBmatcher.cc:1540-1667
int orbm_search_by_projection_keyframe(orbm_t* m, const orbm_frame* cur, const orbm_mappoint* pts, int npts,
                                       const float* minInv, const float* maxInv, const float* maxDistance, float logScaleFactor,
                                       orbm_predict_scale_fn predict, void* user, const uint8_t* alreadyFound,
                                       const float* kfAngle, const uint8_t* curHasPoint, float th, int ORBdist, int checkOri,
                                       int32_t* curAssign, int* nmatchesOut) {
    if (!m || !cur || (!pts && npts) || npts < 0 || !curAssign || !nmatchesOut || (cur->N > 0 && !curHasPoint)) return ORBX_E_INVALID;
    if (npts > 0 && (!minInv || !maxInv || !alreadyFound || !kfAngle || (!predict && !maxDistance))) return ORBX_E_INVALID;
    if (cur->nScaleLevels <= 0) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    DeviceFrame d;
    int rc = upload_frame(m, cur, d);
    if (rc) return rc;
    const float* Tc = cur->Tcw;
    const float tcw[3] = {Tc[3], Tc[7], Tc[11]};
    float Ow[3];                                           // -Rcw.t()*tcw (general gemm path: double accumulation)
    for (int i = 0; i < 3; i++) {
        double s = 0;
        for (int k = 0; k < 3; k++) s += (double)Tc[4 * k + i] * (double)tcw[k];
        Ow[i] = (float)(s * -1.0);
    }
    std::vector<int> meta; std::vector<WQuery> queries;
    for (int i = 0; i < npts; i++) {
        if (!pts[i].valid) continue;
        if (pts[i].bad || alreadyFound[i]) continue;
        float x3Dc[3];
        mat3_mul_add(Tc, 4, pts[i].world, tcw, x3Dc);
        const float xc = x3Dc[0], yc = x3Dc[1];
        const float invzc = (float)(1.0 / x3Dc[2]);
        const float u = cur->fx * xc * invzc + cur->cx;
        const float v = cur->fy * yc * invzc + cur->cy;
        if (u < cur->mnMinX || u > cur->mnMaxX) continue;
        if (v < cur->mnMinY || v > cur->mnMaxY) continue;
        // cv::norm(x3Dw - Ow): float differences, squares summed in double, square root in double (:1581-1582)
        double s2 = 0;
        for (int k = 0; k < 3; k++) { const float po = pts[i].world[k] - Ow[k]; s2 += (double)po * (double)po; }
        const float dist3D = (float)std::sqrt(s2);
        if (dist3D < minInv[i] || dist3D > maxInv[i]) continue;
        int nPredictedLevel;
        if (predict) nPredictedLevel = predict(user, i, dist3D);
        else {                                              // MapPoint::PredictScale, src/MapPoint.cc:402-417
            const float ratio = maxDistance[i] / dist3D;
            nPredictedLevel = (int)std::ceil(std::log(ratio) / logScaleFactor);
            if (nPredictedLevel < 0) nPredictedLevel = 0;
            else if (nPredictedLevel >= cur->nScaleLevels) nPredictedLevel = cur->nScaleLevels - 1;
        }
        if (nPredictedLevel < 0 || nPredictedLevel >= cur->nScaleLevels) return ORBX_E_INVALID;
        WQuery q; std::memset(&q, 0, sizeof(q));
        q.x = u; q.y = v; q.r = th * cur->scaleFactors[nPredictedLevel];
        q.minLevel = nPredictedLevel - 1; q.maxLevel = nPredictedLevel + 1;
        set_desc(q, pts[i].descriptor);
        queries.push_back(q); meta.push_back(i);
    }
    std::vector<int32_t> offs; std::vector<int2> pairs;
    if ((rc = run_queries(m, d, queries, offs, pairs))) return rc;
    int nmatches = 0;
    std::vector<uint8_t> has(curHasPoint, curHasPoint + cur->N);
    for (int i = 0; i < cur->N; i++) curAssign[i] = -2;
    std::vector<int> rotHist[ORBM_HISTO_LENGTH];
    for (size_t qi = 0; qi < queries.size(); qi++) {
        if (offs[qi + 1] == offs[qi]) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (int k = offs[qi]; k < offs[qi + 1]; k++) {
            const int i2 = pairs[k].x;
            if (has[i2]) continue;                                             // :1613-1614
            const int dist = pairs[k].y;
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= ORBdist) {
            curAssign[bestIdx2] = meta[qi]; has[bestIdx2] = 1;
            nmatches++;
            if (checkOri) rotHist[rot_bin(kfAngle[meta[qi]], cur->keysUn[bestIdx2].angle)].push_back(bestIdx2);
        }
    }
    if (checkOri) {
        int32_t counts[ORBM_HISTO_LENGTH];
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) counts[i] = (int)rotHist[i].size();
        int ind1 = -1, ind2 = -1, ind3 = -1;
        orbm_three_maxima(counts, ORBM_HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) { curAssign[idx] = -1; nmatches--; }
    }
    *nmatchesOut = nmatches;
    return ORBX_OK;
}

// ORBmatcher::Fuse(pKF, vpMapPoints, th), src/ORBmatcher.cc:893-1040: selection of the keypoint every map point fuses with
int orbm_fuse_select(orbm_t* m, const orbm_frame* kf, const float* cameraCenter, const float* invLevelSigma2,
                     const orbm_mappoint* pts, int npts,
                     const uint8_t* inKeyFrame, const float* normals, const float* minInv, const float* maxInv, const float* maxDistance,
                     float logScaleFactor, orbm_predict_scale_fn predict, void* user, float th, int32_t* bestIdxOut) {
    if (!m || !kf || !invLevelSigma2 || (!pts && npts) || npts < 0 || !bestIdxOut) return ORBX_E_INVALID;
    if (npts > 0 && (!inKeyFrame || !normals || !minInv || !maxInv || (!predict && !maxDistance))) return ORBX_E_INVALID;
    if (kf->nScaleLevels <= 0 || (kf->N > 0 && !kf->uRight)) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    DeviceFrame d;
    int rc = upload_frame(m, kf, d);
    if (rc) return rc;
    const float* Tc = kf->Tcw;
    const float tcw[3] = {Tc[3], Tc[7], Tc[11]};
    float Ow[3];                                           // GetCameraCenter() = -Rcw.t()*tcw (KeyFrame::SetPose, src/KeyFrame.cc:70-85)
    for (int i = 0; i < 3; i++) {
        double s = 0;
        for (int k = 0; k < 3; k++) s += (double)Tc[4 * k + i] * (double)tcw[k];
        Ow[i] = cameraCenter ? cameraCenter[i] : (float)(s * -1.0);
    }
    struct Q { int i; float u, v, ur; };
    std::vector<Q> meta; std::vector<WQuery> queries;
    for (int i = 0; i < npts; i++) {
        bestIdxOut[i] = -1;
        if (!pts[i].valid) continue;
        if (pts[i].bad || inKeyFrame[i]) continue;
        float p3Dc[3];
        mat3_mul_add(Tc, 4, pts[i].world, tcw, p3Dc);
        if (p3Dc[2] < 0.0f) continue;                                          // depth must be positive
        const float invz = 1 / p3Dc[2];
        const float x = p3Dc[0] * invz, y = p3Dc[1] * invz;
        const float u = kf->fx * x + kf->cx, v = kf->fy * y + kf->cy;
        if (!(u >= kf->mnMinX && u < kf->mnMaxX && v >= kf->mnMinY && v < kf->mnMaxY)) continue;    // KeyFrame::IsInImage
        const float ur = u - kf->bf * invz;
        float PO[3];
        double s2 = 0, dot = 0;
        for (int k = 0; k < 3; k++) {
            PO[k] = pts[i].world[k] - Ow[k];
            s2 += (double)PO[k] * (double)PO[k];                               // cv::norm: double accumulator
            dot += (double)PO[k] * (double)normals[3 * i + k];                 // cv::Mat::dot on CV_32F: double accumulator
        }
        const float dist3D = (float)std::sqrt(s2);
        if (dist3D < minInv[i] || dist3D > maxInv[i]) continue;
        if (dot < 0.5 * dist3D) continue;                                      // viewing angle below 60 degrees
        int nPredictedLevel;
        if (predict) nPredictedLevel = predict(user, i, dist3D);
        else {
            const float ratio = maxDistance[i] / dist3D;
            nPredictedLevel = (int)std::ceil(std::log(ratio) / logScaleFactor);
            if (nPredictedLevel < 0) nPredictedLevel = 0;
            else if (nPredictedLevel >= kf->nScaleLevels) nPredictedLevel = kf->nScaleLevels - 1;
        }
        if (nPredictedLevel < 0 || nPredictedLevel >= kf->nScaleLevels) return ORBX_E_INVALID;
        WQuery q; std::memset(&q, 0, sizeof(q));
        q.x = u; q.y = v; q.r = th * kf->scaleFactors[nPredictedLevel];
        q.minLevel = nPredictedLevel - 1; q.maxLevel = nPredictedLevel;        // the kpLevel gate of :972-973
        set_desc(q, pts[i].descriptor);
        queries.push_back(q); meta.push_back(Q{i, u, v, ur});
    }
    std::vector<int32_t> offs; std::vector<int2> pairs;
    if ((rc = run_queries(m, d, queries, offs, pairs))) return rc;
    for (size_t qi = 0; qi < queries.size(); qi++) {
        const Q& mq = meta[qi];
        int bestDist = 256, bestIdx = -1;
        for (int k = offs[qi]; k < offs[qi + 1]; k++) {
            const int idx = pairs[k].x;
            const orbx_keypoint& kp = kf->keysUn[idx];
            const float ex = mq.u - kp.x, ey = mq.v - kp.y;
            if (kf->uRight[idx] >= 0) {                                        // stereo reprojection error, :975-989
                const float er = mq.ur - kf->uRight[idx];
                const float e2 = ex * ex + ey * ey + er * er;
                if (e2 * invLevelSigma2[kp.octave] > 7.8) continue;
            } else {
                const float e2 = ex * ex + ey * ey;
                if (e2 * invLevelSigma2[kp.octave] > 5.99) continue;
            }
            const int dist = pairs[k].y;
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestDist <= ORBM_TH_LOW) bestIdxOut[mq.i] = bestIdx;
    }
    return ORBX_OK;
}

}  // extern "C"

namespace {
// Projection part shared by SearchByProjection(pKF, Scw, ...) (:318-359) and Fuse(pKF, Scw, ...) (:1072-1113): one window
// query per point that survives the depth, image, distance-range and viewing-angle gates.  invzDouble: "1.0/z" (:1086)
// instead of "1/z" (:327).
int sim3_queries(const orbm_frame* kf, const float* Rcw, const float* tcw, const float* Ow, const orbm_mappoint* pts, int npts,
                 const uint8_t* excluded, const float* normals, const float* minInv, const float* maxInv, const float* maxDistance,
                 float logScaleFactor, orbm_predict_scale_fn predict, void* user, bool invzDouble, float th,
                 std::vector<WQuery>& queries, std::vector<int>& meta) {
    for (int i = 0; i < npts; i++) {
        if (!pts[i].valid) return ORBX_E_INVALID;
        if (excluded[i]) continue;
        float p3Dc[3];
        mat3_mul_add(Rcw, 3, pts[i].world, tcw, p3Dc);
        if (p3Dc[2] < 0.0f) continue;
        const float invz = invzDouble ? (float)(1.0 / p3Dc[2]) : 1 / p3Dc[2];
        const float x = p3Dc[0] * invz, y = p3Dc[1] * invz;
        const float u = kf->fx * x + kf->cx, v = kf->fy * y + kf->cy;
        if (!(u >= kf->mnMinX && u < kf->mnMaxX && v >= kf->mnMinY && v < kf->mnMaxY)) continue;    // KeyFrame::IsInImage
        double s2 = 0, dot = 0;
        for (int k = 0; k < 3; k++) {
            const float po = pts[i].world[k] - Ow[k];
            s2 += (double)po * (double)po;
            dot += (double)po * (double)normals[3 * i + k];
        }
        const float dist = (float)std::sqrt(s2);
        if (dist < minInv[i] || dist > maxInv[i]) continue;
        if (dot < 0.5 * dist) continue;
        int nPredictedLevel;
        if (predict) nPredictedLevel = predict(user, i, dist);
        else {
            const float ratio = maxDistance[i] / dist;
            nPredictedLevel = (int)std::ceil(std::log(ratio) / logScaleFactor);
            if (nPredictedLevel < 0) nPredictedLevel = 0;
            else if (nPredictedLevel >= kf->nScaleLevels) nPredictedLevel = kf->nScaleLevels - 1;
        }
        if (nPredictedLevel < 0 || nPredictedLevel >= kf->nScaleLevels) return ORBX_E_INVALID;
        WQuery q; std::memset(&q, 0, sizeof(q));
        q.x = u; q.y = v; q.r = th * kf->scaleFactors[nPredictedLevel];
        q.minLevel = nPredictedLevel - 1; q.maxLevel = nPredictedLevel;
        set_desc(q, pts[i].descriptor);
        queries.push_back(q); meta.push_back(i);
    }
    return 0;
}
}  // namespace

extern "C" {

int orbm_search_by_projection_sim3(orbm_t* m, const orbm_frame* kf, const float* Rcw, const float* tcw, const float* Ow,
                                   const orbm_mappoint* pts, int npts, const uint8_t* excluded, const float* normals,
                                   const float* minInv, const float* maxInv, const float* maxDistance, float logScaleFactor,
                                   orbm_predict_scale_fn predict, void* user, const uint8_t* matchedIn, int th, int32_t* match,
                                   int* nmatchesOut) {
    if (!m || !kf || !Rcw || !tcw || !Ow || (!pts && npts) || npts < 0 || !match || !nmatchesOut || (kf->N > 0 && !matchedIn)) return ORBX_E_INVALID;
    if (npts > 0 && (!excluded || !normals || !minInv || !maxInv || (!predict && !maxDistance))) return ORBX_E_INVALID;
    if (kf->nScaleLevels <= 0) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    DeviceFrame d;
    int rc = upload_frame(m, kf, d);
    if (rc) return rc;
    std::vector<WQuery> queries; std::vector<int> meta;
    if ((rc = sim3_queries(kf, Rcw, tcw, Ow, pts, npts, excluded, normals, minInv, maxInv, maxDistance, logScaleFactor, predict, user,
                           false, (float)th, queries, meta))) return rc;
    std::vector<int32_t> offs; std::vector<int2> pairs;
    if ((rc = run_queries(m, d, queries, offs, pairs))) return rc;
    std::vector<uint8_t> matched(matchedIn, matchedIn + kf->N);
    for (int i = 0; i < kf->N; i++) match[i] = -2;
    int nmatches = 0;
    for (size_t qi = 0; qi < queries.size(); qi++) {
        int bestDist = 256, bestIdx = -1;
        for (int k = offs[qi]; k < offs[qi + 1]; k++) {
            const int idx = pairs[k].x;
            if (matched[idx]) continue;                                        // :375-376
            const int dist = pairs[k].y;
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestDist <= ORBM_TH_LOW) { match[bestIdx] = meta[qi]; matched[bestIdx] = 1; nmatches++; }
    }
    *nmatchesOut = nmatches;
    return ORBX_OK;
}

int orbm_fuse_select_sim3(orbm_t* m, const orbm_frame* kf, const float* Rcw, const float* tcw, const float* Ow,
                          const orbm_mappoint* pts, int npts, const uint8_t* excluded, const float* normals, const float* minInv,
                          const float* maxInv, const float* maxDistance, float logScaleFactor, orbm_predict_scale_fn predict,
                          void* user, float th, int32_t* bestIdxOut) {
    if (!m || !kf || !Rcw || !tcw || !Ow || (!pts && npts) || npts < 0 || !bestIdxOut) return ORBX_E_INVALID;
    if (npts > 0 && (!excluded || !normals || !minInv || !maxInv || (!predict && !maxDistance))) return ORBX_E_INVALID;
    if (kf->nScaleLevels <= 0) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    DeviceFrame d;
    int rc = upload_frame(m, kf, d);
    if (rc) return rc;
    for (int i = 0; i < npts; i++) bestIdxOut[i] = -1;
    std::vector<WQuery> queries; std::vector<int> meta;
    if ((rc = sim3_queries(kf, Rcw, tcw, Ow, pts, npts, excluded, normals, minInv, maxInv, maxDistance, logScaleFactor, predict, user,
                           true, th, queries, meta))) return rc;
    std::vector<int32_t> offs; std::vector<int2> pairs;
    if ((rc = run_queries(m, d, queries, offs, pairs))) return rc;
    for (size_t qi = 0; qi < queries.size(); qi++) {
        int bestDist = 0x7fffffff, bestIdx = -1;                               // INT_MAX, :1126
        for (int k = offs[qi]; k < offs[qi + 1]; k++)
            if (pairs[k].y < bestDist) { bestDist = pairs[k].y; bestIdx = pairs[k].x; }
        if (bestDist <= ORBM_TH_LOW) bestIdxOut[meta[qi]] = bestIdx;
    }
    return ORBX_OK;
}

}  // extern "C"

namespace {
// one direction of SearchBySim3 (:1215-1294 / :1296-1375): the points of `src` through its own pose and then `sR`,`t`
// into `dst`; best[i] = keypoint of dst with bestDist <= TH_HIGH, else -1
int sim3_direction(orbm_context* m, const orbm_sim3_side* src, const orbm_frame* dst, const float* sR, const float* t, float th,
                   std::vector<int>& best) {
    const orbm_frame* kf = src->kf;
    best.assign(std::max(kf->N, 0), -1);
    DeviceFrame d;
    int rc = upload_frame(m, dst, d);
    if (rc) return rc;
    const float tw[3] = {kf->Tcw[3], kf->Tcw[7], kf->Tcw[11]};
    std::vector<WQuery> queries; std::vector<int> meta;
    for (int i = 0; i < kf->N; i++) {
        if (!src->points[i].valid || src->skip[i]) continue;                     // !pMP || vbAlreadyMatched || isBad()
        float pa[3], pb[3];
        mat3_mul_add(kf->Tcw, 4, src->points[i].world, tw, pa);
        mat3_mul_add(sR, 3, pa, t, pb);
        if (pb[2] < 0.0f) continue;
        const float invz = (float)(1.0 / pb[2]);
        const float x = pb[0] * invz, y = pb[1] * invz;
        const float u = kf->fx * x + kf->cx, v = kf->fy * y + kf->cy;            // pKF1's intrinsics in both directions (:1173-1176)
        if (!(u >= dst->mnMinX && u < dst->mnMaxX && v >= dst->mnMinY && v < dst->mnMaxY)) continue;
        double s2 = 0;
        for (int k = 0; k < 3; k++) s2 += (double)pb[k] * (double)pb[k];
        const float dist3D = (float)std::sqrt(s2);
        if (dist3D < src->min_distance_inv[i] || dist3D > src->max_distance_inv[i]) continue;
        int nPredictedLevel;
        if (src->predict) nPredictedLevel = src->predict(src->user, i, dist3D);
        else {
            const float ratio = src->max_distance[i] / dist3D;
            nPredictedLevel = (int)std::ceil(std::log(ratio) / src->log_scale_factor);
            if (nPredictedLevel < 0) nPredictedLevel = 0;
            else if (nPredictedLevel >= dst->nScaleLevels) nPredictedLevel = dst->nScaleLevels - 1;
        }
        if (nPredictedLevel < 0 || nPredictedLevel >= dst->nScaleLevels) return ORBX_E_INVALID;
        WQuery q; std::memset(&q, 0, sizeof(q));
        q.x = u; q.y = v; q.r = th * dst->scaleFactors[nPredictedLevel];
        q.minLevel = nPredictedLevel - 1; q.maxLevel = nPredictedLevel;
        set_desc(q, src->points[i].descriptor);
        queries.push_back(q); meta.push_back(i);
    }
    std::vector<int32_t> offs; std::vector<int2> pairs;
    if ((rc = run_queries(m, d, queries, offs, pairs))) return rc;
    for (size_t qi = 0; qi < queries.size(); qi++) {
        int bestDist = 0x7fffffff, bestIdx = -1;
        for (int k = offs[qi]; k < offs[qi + 1]; k++)
            if (pairs[k].y < bestDist) { bestDist = pairs[k].y; bestIdx = pairs[k].x; }
        if (bestDist <= ORBM_TH_HIGH) best[meta[qi]] = bestIdx;
    }
    return 0;
}
}  // namespace

extern "C" int orbm_search_by_sim3(orbm_t* m, const orbm_sim3_side* side1, const orbm_sim3_side* side2, const float* sR12, const float* t12,
                                   const float* sR21, const float* t21, float th, int32_t* match12, int* nfound) {
    if (!m || !side1 || !side2 || !side1->kf || !side2->kf || !sR12 || !t12 || !sR21 || !t21 || !match12 || !nfound) return ORBX_E_INVALID;
    const orbm_sim3_side* sd[2] = {side1, side2};
    for (int k = 0; k < 2; k++) {
        const orbm_sim3_side* s = sd[k];
        if (s->kf->N < 0 || s->kf->nScaleLevels <= 0) return ORBX_E_INVALID;
        if (s->kf->N > 0 && (!s->points || !s->skip || !s->min_distance_inv || !s->max_distance_inv || (!s->predict && !s->max_distance)))
            return ORBX_E_INVALID;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    // the intrinsics of pKF1 project in both directions (:1173-1176): direction 2->1 reads them from side1's frame
    orbm_frame k2as1 = *side2->kf;
    k2as1.fx = side1->kf->fx; k2as1.fy = side1->kf->fy; k2as1.cx = side1->kf->cx; k2as1.cy = side1->kf->cy;
    orbm_sim3_side s2 = *side2; s2.kf = &k2as1;
    std::vector<int> vnMatch1, vnMatch2;
    int rc;
    if ((rc = sim3_direction(m, side1, side2->kf, sR21, t21, th, vnMatch1))) return rc;
    if ((rc = sim3_direction(m, &s2, side1->kf, sR12, t12, th, vnMatch2))) return rc;
    int n = 0;
    for (int i1 = 0; i1 < side1->kf->N; i1++) {                                    // :1377-1391
        const int idx2 = vnMatch1[i1];
        match12[i1] = -1;
        if (idx2 >= 0 && vnMatch2[idx2] == i1) { match12[i1] = idx2; n++; }
    }
    *nfound = n;
    return ORBX_OK;
}

// SearchForInitialization over many pairs of device-resident frames (initbatch.cuh)
extern "C" int orbm_search_for_initialization_batch_device(orbm_t* m, const orbx_keypoint* d_keys, const uint8_t* d_desc, const int32_t* d_n,
                                                           int cap, int nframes, const int32_t* d_pairs, int npairs, float mnMinX,
                                                           float mnMinY, float invW, float invH, float* d_prev, int windowSize,
                                                           float nnratio, int checkOri, int32_t* d_matches12, int32_t* d_nmatches) {
    if (!m || !d_keys || !d_desc || !d_n || cap <= 0 || nframes <= 0 || npairs < 0 || !d_matches12 || !d_nmatches) return ORBX_E_INVALID;
    if (!d_pairs && npairs > nframes - 1) return ORBX_E_INVALID;
    if (((uintptr_t)d_desc & 15) || cap >= (1 << 20)) return ORBX_E_INVALID;
    if (npairs == 0) return ORBX_OK;
    std::lock_guard<std::mutex> lk(m->mtx);
    CKM(cudaSetDevice(m->device));
    int rc;
    const size_t fc = (size_t)nframes * cap, pc = (size_t)npairs * cap;
    if ((rc = m->w0.reserve(fc * 4))) return rc;                                   // cellOf
    if ((rc = m->w1.reserve((size_t)nframes * (GRID_CELLS + 1) * 4))) return rc;    // cellStart
    if ((rc = m->w2.reserve(fc * 4))) return rc;                                   // items
    if ((rc = m->w3.reserve(pc * 4))) return rc;                                   // vMatchedDistance
    if ((rc = m->w4.reserve(pc * 4))) return rc;                                   // vnMatches21
    if ((rc = m->w5.reserve(pc))) return rc;                                       // rotation bins
    const GridParams g{mnMinX, mnMinY, invW, invH};
    grid_build_batch_kernel<<<nframes, 1024, 0, m->stream>>>(d_keys, d_n, cap, g, 0, (int32_t*)m->w0.p, (int32_t*)m->w1.p, (int32_t*)m->w2.p);
    InitBatchArgs a{};
    a.keys = d_keys; a.desc = (const uint32_t*)d_desc; a.counts = d_n; a.cap = cap; a.pairs = d_pairs; a.npairs = npairs;
    a.cellStart = (const int32_t*)m->w1.p; a.items = (const int32_t*)m->w2.p; a.g = g;
    a.windowSize = (float)windowSize; a.nnratio = nnratio; a.checkOri = checkOri; a.prev = d_prev;
    a.matchedDistance = (int32_t*)m->w3.p; a.matches21 = (int32_t*)m->w4.p; a.bins = (uint8_t*)m->w5.p;
    a.matches12 = d_matches12; a.nmatches = d_nmatches;
    static const bool noTeam = std::getenv("ORBM_NO_TEAM") != nullptr;
    if (cap <= INIT_TEAM_MAX_CAP && !noTeam) {               // one block per pair: enumerator warps + one resolver warp
        const size_t smem = init_team_smem(cap);
        CKM(cudaFuncSetAttribute(init_match_team_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        init_match_team_kernel<<<npairs, (INIT_TEAM + 1) * 32, smem, m->stream>>>(a);
    } else {
        init_match_batch_kernel<<<(npairs + 3) / 4, 128, 0, m->stream>>>(a);
    }
    CKM(cudaGetLastError());
    return ORBX_OK;
}
