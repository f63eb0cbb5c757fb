// ORACLE — test infrastructure only (see orb_oracle.h).
//
// CPU restatement of
//   ORBVocabulary = DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>  (include/ORBVocabulary.h:30-31):
//     loadFromTextFile                           Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1426
//     transform(features, v, fv, levelsup)       :1126-1194   (Frame::ComputeBoW src/Frame.cc:521-528,
//                                                              KeyFrame::ComputeBoW src/KeyFrame.cc:59-68, levelsup = 4)
//     transform(feature, id, weight, nid, lup)   :1217-1259
//     BowVector::addWeight/addIfNotExist/normalize   Thirdparty/DBoW2/DBoW2/BowVector.cpp:33-89
//     FeatureVector::addFeature                      Thirdparty/DBoW2/DBoW2/FeatureVector.cpp:27-42
//     FORB::distance                                 Thirdparty/DBoW2/DBoW2/FORB.cpp:84-104
//   MapPoint::ComputeDistinctiveDescriptors      src/MapPoint.cc:241-306 (the distance matrix + median selection)
// Pinned: the transform is bit-exact against the reference's own DBoW2 sources compiled in oracle/_ref
// (libdbow2_ref.so, `make ref`; tests/test_oracle_vocab_cpu.py); committed golden vectors for the GPU box
// (tests/golden/vocab_*.npz) come from that build.
// Canonical choice: loadFromTextFile's blank-line artefact (an extra root child with an uninitialised descriptor
// when the file ends in a newline) is not reproduced - blank lines are skipped.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

namespace {

typedef unsigned int WordId;
typedef unsigned int NodeId;
typedef double WordValue;
enum LNorm { L1, L2 };
enum WeightingType { TF_IDF, TF, IDF, BINARY };
enum ScoringType { L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT };

struct Node {                                        // TemplatedVocabulary.h:297-329
    NodeId id; WordValue weight; std::vector<NodeId> children; NodeId parent; uint8_t descriptor[32]; WordId word_id;
    Node() : id(0), weight(0), parent(0), word_id(0) { std::memset(descriptor, 0, 32); }
    bool isLeaf() const { return children.empty(); }
};

struct Vocabulary {
    int m_k, m_L; WeightingType m_weighting; ScoringType m_scoring;
    std::vector<Node> m_nodes; std::vector<NodeId> m_words;          // m_words holds node ids (pointers in the reference)
    bool empty() const { return m_words.empty(); }
};

int distance(const uint8_t* a, const uint8_t* b) {   // FORB.cpp:84-104
    int32_t pa[8], pb[8];
    std::memcpy(pa, a, 32); std::memcpy(pb, b, 32);
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        unsigned int v = pa[i] ^ pb[i];
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

bool mustNormalize(ScoringType s, LNorm& norm) {     // ScoringObject.h:74-89
    norm = s == L2_NORM ? L2 : L1;
    return s != DOT_PRODUCT;
}

void transform1(const Vocabulary& V, const uint8_t* feature, WordId& word_id, WordValue& weight, NodeId* nid, int levelsup) {
    std::vector<NodeId> nodes;                       // :1221-1258
    const int nid_level = V.m_L - levelsup;
    if (nid_level <= 0 && nid != NULL) *nid = 0;
    NodeId final_id = 0;
    int current_level = 0;
    do {
        ++current_level;
        nodes = V.m_nodes[final_id].children;
        final_id = nodes[0];
        double best_d = distance(feature, V.m_nodes[final_id].descriptor);
        for (std::vector<NodeId>::const_iterator nit = nodes.begin() + 1; nit != nodes.end(); ++nit) {
            NodeId id = *nit;
            double d = distance(feature, V.m_nodes[id].descriptor);
            if (d < best_d) { best_d = d; final_id = id; }
        }
        if (nid != NULL && current_level == nid_level) *nid = final_id;
    } while (!V.m_nodes[final_id].isLeaf());
    word_id = V.m_nodes[final_id].word_id;
    weight = V.m_nodes[final_id].weight;
}

typedef std::map<WordId, WordValue> BowVector;
typedef std::map<NodeId, std::vector<unsigned int> > FeatureVector;

void addWeight(BowVector& v, WordId id, WordValue w) {           // BowVector.cpp:33-45
    BowVector::iterator vit = v.lower_bound(id);
    if (vit != v.end() && !(v.key_comp()(id, vit->first))) vit->second += w;
    else v.insert(vit, BowVector::value_type(id, w));
}
void addIfNotExist(BowVector& v, WordId id, WordValue w) {       // BowVector.cpp:49-57
    BowVector::iterator vit = v.lower_bound(id);
    if (vit == v.end() || (v.key_comp()(id, vit->first))) v.insert(vit, BowVector::value_type(id, w));
}
void normalize(BowVector& v, LNorm norm_type) {                  // BowVector.cpp:61-83
    double norm = 0.0;
    if (norm_type == L1) { for (BowVector::iterator it = v.begin(); it != v.end(); ++it) norm += std::fabs(it->second); }
    else { for (BowVector::iterator it = v.begin(); it != v.end(); ++it) norm += it->second * it->second; norm = std::sqrt(norm); }
    if (norm > 0.0) for (BowVector::iterator it = v.begin(); it != v.end(); ++it) it->second /= norm;
}
void addFeature(FeatureVector& fv, NodeId id, unsigned int i_feature) {   // FeatureVector.cpp:27-42
    FeatureVector::iterator vit = fv.lower_bound(id);
    if (vit != fv.end() && vit->first == id) vit->second.push_back(i_feature);
    else { vit = fv.insert(vit, FeatureVector::value_type(id, std::vector<unsigned int>())); vit->second.push_back(i_feature); }
}

void transform(const Vocabulary& V, const uint8_t* features, int n, BowVector& v, FeatureVector& fv, int levelsup) {   // :1126-1194
    v.clear(); fv.clear();
    if (V.empty()) return;
    LNorm norm;
    bool must = mustNormalize(V.m_scoring, norm);
    if (V.m_weighting == TF || V.m_weighting == TF_IDF) {
        for (unsigned int i_feature = 0; i_feature < (unsigned)n; ++i_feature) {
            WordId id; NodeId nid = 0; WordValue w;
            transform1(V, features + (size_t)i_feature * 32, id, w, &nid, levelsup);
            if (w > 0) { addWeight(v, id, w); addFeature(fv, nid, i_feature); }
        }
        if (!v.empty() && !must) {
            const double nd = v.size();
            for (BowVector::iterator vit = v.begin(); vit != v.end(); vit++) vit->second /= nd;
        }
    } else {
        for (unsigned int i_feature = 0; i_feature < (unsigned)n; ++i_feature) {
            WordId id; NodeId nid = 0; WordValue w;
            transform1(V, features + (size_t)i_feature * 32, id, w, &nid, levelsup);
            if (w > 0) { addIfNotExist(v, id, w); addFeature(fv, nid, i_feature); }
        }
    }
    if (must) normalize(v, norm);
}

}  // namespace

extern "C" {

void* orbo_vocab_create(int k, int L, int scoring, int weighting, int nnodes, const int32_t* parent, const uint8_t* is_leaf,
                        const uint8_t* desc, const double* weight) {
    Vocabulary* V = new Vocabulary();
    V->m_k = k; V->m_L = L; V->m_scoring = (ScoringType)scoring; V->m_weighting = (WeightingType)weighting;
    V->m_nodes.resize(1);
    V->m_nodes[0].id = 0;
    for (int nid = 1; nid < nnodes; nid++) {         // the node loop of loadFromTextFile, :1379-1421
        V->m_nodes.resize(V->m_nodes.size() + 1);
        V->m_nodes[nid].id = nid;
        const int pid = parent[nid];
        V->m_nodes[nid].parent = pid;
        V->m_nodes[pid].children.push_back(nid);
        std::memcpy(V->m_nodes[nid].descriptor, desc + (size_t)nid * 32, 32);
        V->m_nodes[nid].weight = weight[nid];
        if (is_leaf[nid] > 0) {
            const int wid = (int)V->m_words.size();
            V->m_words.resize(wid + 1);
            V->m_nodes[nid].word_id = wid;
            V->m_words[wid] = nid;
        }
    }
    return V;
}

void* orbo_vocab_load_text(const char* filename) {   // :1338-1426, iostream parsing as in the reference
    std::ifstream f(filename);
    if (!f.is_open()) return NULL;
    std::string s;
    std::getline(f, s);
    std::stringstream ss;
    ss << s;
    int k, L, n1, n2;
    ss >> k; ss >> L; ss >> n1; ss >> n2;
    if (k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) return NULL;
    std::vector<int32_t> parent(1, 0);
    std::vector<uint8_t> leaf(1, 0), desc(32, 0);
    std::vector<double> weight(1, 0.0);
    while (!f.eof()) {
        std::string snode;
        std::getline(f, snode);
        if (snode.find_first_not_of(" \t\r\n") == std::string::npos) continue;       // canonical: skip blank lines
        std::stringstream ssnode;
        ssnode << snode;
        int pid; ssnode >> pid;
        int nIsLeaf; ssnode >> nIsLeaf;
        uint8_t d[32];
        for (int iD = 0; iD < 32; iD++) { int e; ssnode >> e; d[iD] = (unsigned char)e; }      // FORB::fromString, FORB.cpp:122-137
        double w; ssnode >> w;
        parent.push_back(pid); leaf.push_back(nIsLeaf > 0); desc.insert(desc.end(), d, d + 32); weight.push_back(w);
    }
    return orbo_vocab_create(k, L, n1, n2, (int)parent.size(), parent.data(), leaf.data(), desc.data(), weight.data());
}

void orbo_vocab_destroy(void* h) { delete (Vocabulary*)h; }
int orbo_vocab_words(void* h) { return (int)((Vocabulary*)h)->m_words.size(); }
int orbo_vocab_nodes(void* h) { return (int)((Vocabulary*)h)->m_nodes.size(); }

void orbo_vocab_descend(void* h, const uint8_t* desc, int n, int levelsup, uint32_t* word_id, uint32_t* node_id, double* weight) {
    const Vocabulary& V = *(Vocabulary*)h;
    for (int i = 0; i < n; i++) {
        WordId id; NodeId nid = 0; WordValue w;
        transform1(V, desc + (size_t)i * 32, id, w, &nid, levelsup);
        word_id[i] = id; node_id[i] = nid; weight[i] = w;
    }
}

int orbo_vocab_transform(void* h, const uint8_t* desc, int n, int levelsup, uint32_t* word_ids, double* word_vals, int* nwords,
                         uint32_t* node_ids, int32_t* node_start, uint32_t* items, int* nnodes) {
    BowVector bow; FeatureVector fv;
    transform(*(Vocabulary*)h, desc, n, bow, fv, levelsup);
    int k = 0;
    for (BowVector::const_iterator it = bow.begin(); it != bow.end(); ++it, ++k) { word_ids[k] = it->first; word_vals[k] = it->second; }
    *nwords = k;
    int nn = 0, pos = 0;
    for (FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++nn) {
        node_ids[nn] = it->first; node_start[nn] = pos;
        for (size_t j = 0; j < it->second.size(); j++) items[pos++] = it->second[j];
    }
    node_start[nn] = pos;
    *nnodes = nn;
    return 0;
}

// MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:274-301, for each point's descriptor list
void orbo_distinctive_descriptors(const uint8_t* desc, const int32_t* start, int npoints, int32_t* best) {
    for (int p = 0; p < npoints; p++) {
        const uint8_t* vDescriptors = desc + (size_t)start[p] * 32;
        const size_t N = (size_t)(start[p + 1] - start[p]);
        if (N == 0) { best[p] = -1; continue; }      // vDescriptors.empty() -> return (:272-273)
        std::vector<float> Distances(N * N);
        for (size_t i = 0; i < N; i++) {
            Distances[i * N + i] = 0;
            for (size_t j = i + 1; j < N; j++) {
                int distij = distance(vDescriptors + i * 32, vDescriptors + j * 32);   // ORBmatcher::DescriptorDistance == FORB::distance
                Distances[i * N + j] = distij;
                Distances[j * N + i] = distij;
            }
        }
        int BestMedian = INT_MAX;
        int BestIdx = 0;
        for (size_t i = 0; i < N; i++) {
            std::vector<int> vDists(Distances.begin() + i * N, Distances.begin() + i * N + N);
            std::sort(vDists.begin(), vDists.end());
            int median = vDists[0.5 * (N - 1)];
            if (median < BestMedian) { BestMedian = median; BestIdx = i; }
        }
        best[p] = BestIdx;
    }
}

}  // extern "C"
