// oracle/bow.cpp — CPU ORACLE (TEST INFRASTRUCTURE) for SURVEY.md §8f N2:
//   Frame::ComputeBoW                               /root/reference/src/Frame.cc:447-454  (transform(..., levelsup = 4))
//   TemplatedVocabulary::transform (features)       /root/reference/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1140-1207
//   TemplatedVocabulary::transform (one feature)    .../TemplatedVocabulary.h:1231-1272
//   BowVector::addWeight / normalize                /root/reference/Thirdparty/DBoW2/DBoW2/BowVector.cpp
//   FeatureVector::addFeature                       /root/reference/Thirdparty/DBoW2/DBoW2/FeatureVector.cpp
//   FORB::distance                                  /root/reference/Thirdparty/DBoW2/DBoW2/FORB.cpp:81-101
//   vocabulary layout = loadFromTextFile            .../TemplatedVocabulary.h:1351-1436 (node id = line number, children in
//                                                   push_back order, word ids in order of appearance of the leaves)
// DBoW2 is vendored in the reference but needs OpenCV headers (cv::Mat descriptors), so it cannot be compiled here.
// Only the configuration ORB-SLAM2 ships is restated: weighting TF_IDF, scoring L1_NORM (first line of ORBvoc.txt:
// "10 6 0 0"). The vocabulary file itself is not in the reference checkout; tests use seeded synthetic trees.
#include <cmath>
#include <cstring>
#include <map>
#include <vector>

#include "orb_oracle.h"

struct orc_vocabulary {
    int L = 0;
    struct Node {
        int parent = 0, word_id = 0;
        double weight = 0;
        uint8_t desc[32] = {0};
        std::vector<int> children;
    };
    std::vector<Node> nodes;  // [0] = root
    int n_words = 0;
};

extern "C" orc_vocabulary* orc_voc_create(int L, int n_nodes, const int* parent, const uint8_t* is_leaf, const uint8_t* desc, const double* weight)
{
    orc_vocabulary* v = new orc_vocabulary();
    v->L = L;
    v->nodes.resize(1);
    for (int i = 0; i < n_nodes; ++i) {  // loadFromTextFile :1391-1433
        const int nid = (int)v->nodes.size();
        v->nodes.resize(nid + 1);
        const int pid = parent[i];
        if (pid < 0 || pid >= nid) { delete v; return nullptr; }
        v->nodes[nid].parent = pid;
        v->nodes[pid].children.push_back(nid);
        std::memcpy(v->nodes[nid].desc, desc + 32 * (size_t)i, 32);
        v->nodes[nid].weight = weight[i];
        if (is_leaf[i]) v->nodes[nid].word_id = v->n_words++;
    }
    return v;
}

extern "C" void orc_voc_destroy(orc_vocabulary* v) { delete v; }
extern "C" int orc_voc_words(const orc_vocabulary* v) { return v->n_words; }

static int forb_distance(const uint8_t* a, const uint8_t* b)
{
    int dist = 0;
    for (int i = 0; i < 8; i++) {  // FORB.cpp:90-98
        uint32_t pa, pb;
        std::memcpy(&pa, a + 4 * i, 4);
        std::memcpy(&pb, b + 4 * i, 4);
        unsigned int v = pa ^ pb;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

// transform(feature, word_id, weight, nid, levelsup) :1231-1272. nid starts at 0 (the reference leaves it
// uninitialised when the descent ends above nid_level; balanced vocabularies never get there).
static void transform_one(const orc_vocabulary* voc, const uint8_t* feature, int& word_id, double& weight, int& nid, int levelsup)
{
    const int nid_level = voc->L - levelsup;
    nid = 0;
    int final_id = 0, current_level = 0;
    do {
        ++current_level;
        const std::vector<int>& nodes = voc->nodes[final_id].children;
        final_id = nodes[0];
        double best_d = forb_distance(feature, voc->nodes[final_id].desc);
        for (size_t k = 1; k < nodes.size(); ++k) {
            const int id = nodes[k];
            const double d = forb_distance(feature, voc->nodes[id].desc);
            if (d < best_d) { best_d = d; final_id = id; }
        }
        if (current_level == nid_level) nid = final_id;
    } while (!voc->nodes[final_id].children.empty());
    word_id = voc->nodes[final_id].word_id;
    weight = voc->nodes[final_id].weight;
}

extern "C" int orc_compute_bow(const orc_vocabulary* voc, const uint8_t* desc, int n, int levelsup, int* word_ids, double* word_vals,
                               int* n_words, int* fv_nodes, int* fv_start, int* fv_feats, int* n_fv_nodes)
{
    std::map<int, double> bow;
    std::map<int, std::vector<unsigned>> fv;
    if (voc->nodes.size() > 1 && !voc->nodes[0].children.empty()) {
        for (int i = 0; i < n; ++i) {  // :1160-1175, TF_IDF
            int id, nid;
            double w;
            transform_one(voc, desc + 32 * (size_t)i, id, w, nid, levelsup);
            if (w > 0) {
                auto it = bow.lower_bound(id);  // BowVector::addWeight
                if (it != bow.end() && it->first == id) it->second += w;
                else bow.insert(it, std::make_pair(id, w));
                fv[nid].push_back((unsigned)i);  // FeatureVector::addFeature
            }
        }
        double norm = 0.0;  // BowVector::normalize(L1) — L1_NORM scoring must normalise (:1206)
        for (auto& kv : bow) norm += std::fabs(kv.second);
        if (norm > 0.0)
            for (auto& kv : bow) kv.second /= norm;
    }
    int k = 0;
    for (auto& kv : bow) { word_ids[k] = kv.first; word_vals[k] = kv.second; ++k; }
    *n_words = k;
    int m = 0, f = 0;
    for (auto& kv : fv) {
        fv_nodes[m] = kv.first;
        fv_start[m] = f;
        for (unsigned idx : kv.second) fv_feats[f++] = (int)idx;
        ++m;
    }
    fv_start[m] = f;
    *n_fv_nodes = m;
    return 0;
}
