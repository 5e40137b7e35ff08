// oracle/distribute.cpp — CPU restatement of ORBextractor::DistributeOctTree and
// ExtractorNode::DivideNode (/root/reference/src/ORBextractor.cc:484-540, 542-766).
// TEST INFRASTRUCTURE (see orb_oracle.h).
//
// The list discipline of the reference is kept literally (std::list, children pushed to the front in
// the order UL, UR, BL, BR, parent erased) because the output order of a level IS the final list
// order (:747-763). One thing is canonicalised: the reference sorts (size, ExtractorNode*) pairs
// (:684-688), i.e. equal sizes are ordered by heap address. Here the second key is the node's
// creation sequence number, so equal sizes are expanded later-created-first.
#include <algorithm>
#include <cmath>
#include <list>
#include <utility>
#include <vector>

#include "oracle_internal.h"

namespace {

struct Node {
    int x0, y0, x1, y1;  // UL.x, UL.y, UR.x, BL.y
    std::vector<int> keys;  // indices into the candidate array, input order preserved
    bool leaf = false;      // bNoMore
    long seq = 0;
    std::list<Node>::iterator self;
};

// DivideNode :484-540. Children order: 0 = UL (n1), 1 = UR (n2), 2 = BL (n3), 3 = BR (n4).
void divide(const Node& p, const std::vector<orc_keypoint>& cand, Node child[4])
{
    const int halfX = (int)std::ceil((float)(p.x1 - p.x0) / 2);
    const int halfY = (int)std::ceil((float)(p.y1 - p.y0) / 2);
    const int mx = p.x0 + halfX, my = p.y0 + halfY;
    child[0].x0 = p.x0; child[0].x1 = mx;   child[0].y0 = p.y0; child[0].y1 = my;
    child[1].x0 = mx;   child[1].x1 = p.x1; child[1].y0 = p.y0; child[1].y1 = my;
    child[2].x0 = p.x0; child[2].x1 = mx;   child[2].y0 = my;   child[2].y1 = p.y1;
    child[3].x0 = mx;   child[3].x1 = p.x1; child[3].y0 = my;   child[3].y1 = p.y1;
    for (int k : p.keys) {
        const orc_keypoint& kp = cand[k];
        int q = (kp.x < (float)mx ? 0 : 1) + (kp.y < (float)my ? 0 : 2);
        child[q].keys.push_back(k);
    }
    for (int q = 0; q < 4; ++q) child[q].leaf = child[q].keys.size() == 1;
}

}  // namespace

void orc_distribute_impl(const std::vector<orc_keypoint>& cand, int minX, int maxX, int minY, int maxY, int N,
                         std::vector<orc_keypoint>& out)
{
    out.clear();
    const int nIni = (int)std::round((float)(maxX - minX) / (maxY - minY));  // :546
    if (nIni < 1) return;
    const float hX = (float)(maxX - minX) / nIni;  // :548
    long seq = 0;
    std::list<Node> nodes;
    std::vector<Node*> ini(nIni);
    for (int i = 0; i < nIni; ++i) {  // :555-568
        Node n;
        n.x0 = (int)(hX * (float)i);
        n.x1 = (int)(hX * (float)(i + 1));
        n.y0 = 0;
        n.y1 = maxY - minY;
        n.seq = seq++;
        nodes.push_back(n);
        ini[i] = &nodes.back();
    }
    for (int k = 0; k < (int)cand.size(); ++k) {  // :571-575
        int b = (int)(cand[k].x / hX);
        if (b >= nIni) b = nIni - 1;  // cannot happen for x < width; guards the reference's unchecked index
        ini[b]->keys.push_back(k);
    }
    for (auto it = nodes.begin(); it != nodes.end();) {  // :577-588
        if (it->keys.size() == 1) { it->leaf = true; ++it; }
        else if (it->keys.empty()) it = nodes.erase(it);
        else ++it;
    }

    std::vector<std::pair<std::pair<int, long>, Node*>> toExpand;  // ((size, seq), node)
    // push the non-empty children of a divided node to the list front, n1..n4 (:624-663 / :696-727)
    auto push_children = [&](Node child[4]) {
        for (int q = 0; q < 4; ++q) {
            if (child[q].keys.empty()) continue;
            child[q].seq = seq++;
            nodes.push_front(child[q]);
            if (child[q].keys.size() > 1) {
                nodes.front().self = nodes.begin();
                toExpand.push_back({{(int)child[q].keys.size(), nodes.front().seq}, &nodes.front()});
            }
        }
    };

    bool finish = false;
    while (!finish) {  // :599
        const int prev = (int)nodes.size();
        toExpand.clear();
        for (auto it = nodes.begin(); it != nodes.end();) {  // phase-1 pass :609-668
            if (it->leaf) { ++it; continue; }
            Node child[4];
            divide(*it, cand, child);
            push_children(child);
            it = nodes.erase(it);
        }
        const int nToExpand = (int)toExpand.size();
        if ((int)nodes.size() >= N || (int)nodes.size() == prev) {  // :672
            finish = true;
        } else if ((int)nodes.size() + nToExpand * 3 > N) {  // :676 phase 2
            while (!finish) {
                const int prev2 = (int)nodes.size();
                auto order = toExpand;
                toExpand.clear();
                std::sort(order.begin(), order.end(),
                          [](const auto& a, const auto& b) { return a.first < b.first; });  // (size, seq) ascending
                for (int j = (int)order.size() - 1; j >= 0; --j) {  // largest first :688
                    Node* p = order[j].second;
                    Node child[4];
                    divide(*p, cand, child);
                    push_children(child);
                    nodes.erase(p->self);
                    if ((int)nodes.size() >= N) break;  // :733
                }
                if ((int)nodes.size() >= N || (int)nodes.size() == prev2) finish = true;  // :737
            }
        }
    }

    out.reserve(nodes.size());
    for (const Node& n : nodes) {  // :744-763 best response per node, first wins ties
        int best = n.keys[0];
        float bestR = cand[best].response;
        for (size_t k = 1; k < n.keys.size(); ++k)
            if (cand[n.keys[k]].response > bestR) { best = n.keys[k]; bestR = cand[best].response; }
        out.push_back(cand[best]);
    }
}

extern "C" int orc_distribute(const orc_keypoint* cand, int n, int min_x, int max_x, int min_y, int max_y,
                              int n_target, orc_keypoint* out, int capacity)
{
    std::vector<orc_keypoint> c(cand, cand + n), o;
    orc_distribute_impl(c, min_x, max_x, min_y, max_y, n_target, o);
    for (int i = 0; i < (int)o.size() && i < capacity; ++i) out[i] = o[i];
    return (int)o.size();
}
