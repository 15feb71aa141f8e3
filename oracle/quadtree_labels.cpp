// K3 (host form) — quadtree keypoint distribution, the replacement of
// ORBextractor::DistributeOctTree + ExtractorNode::divideNode
// (Features/ORBextractor.cpp:414-611, Features/ExtractorNode.cpp:5-57).
//
// Formulation: label-based and pass-synchronous instead of the reference's
// std::list surgery.  Points never move; every point carries the slot id of
// the node that owns it.  A pass (a) counts, for every expandable node, how
// many of its points fall in each quadrant, (b) decides which nodes split
// (all of them in the main phase; in the final phase the largest-first prefix
// that reaches N nodes), (c) relabels the points of split nodes and rebuilds
// the ordered node list with one exclusive scan.  Because stable partitions
// preserve candidate order inside every node, "first key with the maximum
// response" (ORBextractor.cpp:592-608) equals "maximum response, lowest
// candidate index", so no per-node ordering has to be maintained.
// Each step is a data-parallel map / scan / sort: the same code shape runs as
// one CTA per (frame, level) on the device (quadtree.cu).
// Tie rule for the size sort: creation order, later-created = greater
// (the reference compares node addresses there).
#include <vector>
// ORACLE — TEST INFRASTRUCTURE ONLY: CPU form of the label-based quadtree formulation that
// k_quadtree implements, kept to cross-check the formulation against the std::list restatement
// (oracle/orb.cpp) without a GPU.  Packed candidate: x | y << 12 | score << 24.
static inline int cand_x(unsigned c) { return c & 0xfff; }
static inline int cand_y(unsigned c) { return (c >> 12) & 0xfff; }
static inline int cand_s(unsigned c) { return c >> 24; }

#include <algorithm>
#include <cmath>

namespace orc {

namespace {
    struct QNode {
        int ulx, uly, brx, bry;
        int count;
        int seq;
        int child[4]; // tentative quadrant counts
    };
}

void host_quadtree(const unsigned* cand, int n, int minX, int maxX, int minY, int maxY, int N, std::vector<unsigned>& out)
{
    out.clear();
    if (n <= 0)
        return;
    const int nIni = (int)roundf(static_cast<float>(maxX - minX) / (maxY - minY));
    const float hX = static_cast<float>(maxX - minX) / nIni;

    std::vector<QNode> nodes;
    std::vector<int> label(n);
    std::vector<int> list; // ordered node ids (front first)
    int seq = 0;
    for (int i = 0; i < nIni; i++) {
        QNode q;
        q.ulx = (int)(hX * static_cast<float>(i));
        q.brx = (int)(hX * static_cast<float>(i + 1));
        q.uly = 0;
        q.bry = maxY - minY;
        q.count = 0;
        q.seq = seq++;
        nodes.push_back(q);
    }
    for (int p = 0; p < n; p++) {
        int k = (int)((float)cand_x(cand[p]) / hX);
        label[p] = k;
        nodes[k].count++;
    }
    for (int i = 0; i < nIni; i++)
        if (nodes[i].count > 0)
            list.push_back(i);

    auto count_children = [&](const std::vector<int>& expand) {
        for (int id : expand)
            nodes[id].child[0] = nodes[id].child[1] = nodes[id].child[2] = nodes[id].child[3] = 0;
        for (int p = 0; p < n; p++) {
            QNode& q = nodes[label[p]];
            if (q.count <= 1)
                continue;
            const int halfX = (int)ceilf(static_cast<float>(q.brx - q.ulx) / 2);
            const int halfY = (int)ceilf(static_cast<float>(q.bry - q.uly) / 2);
            int right = !(cand_x(cand[p]) < q.ulx + halfX);
            int down = !(cand_y(cand[p]) < q.uly + halfY);
            q.child[right + 2 * down]++; // n1=0 n2=1 n3=2 n4=3
        }
    };
    // split node `id`: allocate children in n1..n4 order, return their ids (nonempty only)
    std::vector<int> remap; // parent id -> first child slot (4 consecutive, -1 when empty)
    auto split = [&](int id, int childIds[4]) {
        const QNode q = nodes[id];
        const int halfX = (int)ceilf(static_cast<float>(q.brx - q.ulx) / 2);
        const int halfY = (int)ceilf(static_cast<float>(q.bry - q.uly) / 2);
        for (int k = 0; k < 4; k++) {
            childIds[k] = -1;
            if (q.child[k] == 0)
                continue;
            QNode c;
            c.ulx = (k & 1) ? q.ulx + halfX : q.ulx;
            c.brx = (k & 1) ? q.brx : q.ulx + halfX;
            c.uly = (k & 2) ? q.uly + halfY : q.uly;
            c.bry = (k & 2) ? q.bry : q.uly + halfY;
            c.count = q.child[k];
            c.seq = seq++;
            childIds[k] = (int)nodes.size();
            nodes.push_back(c);
        }
    };
    auto relabel = [&](const std::vector<int>& childOf /* node id -> 4 child ids, flat */) {
        for (int p = 0; p < n; p++) {
            int id = label[p];
            if ((size_t)id * 4 + 3 >= childOf.size() || childOf[id * 4] == -2)
                continue;
            const QNode& q = nodes[id];
            const int halfX = (int)ceilf(static_cast<float>(q.brx - q.ulx) / 2);
            const int halfY = (int)ceilf(static_cast<float>(q.bry - q.uly) / 2);
            int right = !(cand_x(cand[p]) < q.ulx + halfX);
            int down = !(cand_y(cand[p]) < q.uly + halfY);
            label[p] = childOf[id * 4 + right + 2 * down];
        }
    };

    bool finish = false;
    std::vector<int> expandable; // nodes with > 1 point created by the last pass
    while (!finish) {
        const int prevSize = (int)list.size();
        // main-phase pass: every expandable node in list order
        std::vector<int> expand;
        for (int id : list)
            if (nodes[id].count > 1)
                expand.push_back(id);
        count_children(expand);
        std::vector<int> childOf(nodes.size() * 4, -2);
        std::vector<int> front; // new nodes, in creation order
        expandable.clear();
        int nToExpand = 0;
        for (int id : expand) {
            int c[4];
            split(id, c);
            for (int k = 0; k < 4; k++) {
                childOf[id * 4 + k] = c[k];
                if (c[k] >= 0) {
                    front.push_back(c[k]);
                    if (nodes[c[k]].count > 1) {
                        nToExpand++;
                        expandable.push_back(c[k]);
                    }
                }
            }
        }
        relabel(childOf);
        std::vector<int> newList(front.rbegin(), front.rend()); // push_front order
        for (int id : list)
            if (nodes[id].count == 1)
                newList.push_back(id);
        list.swap(newList);

        if ((int)list.size() >= N || (int)list.size() == prevSize) {
            finish = true;
        } else if (((int)list.size() + nToExpand * 3) > N) {
            while (!finish) {
                const int prev = (int)list.size();
                std::vector<int> order = expandable; // sort ascending (size, seq); process from the back
                std::sort(order.begin(), order.end(), [&](int a, int b) {
                    if (nodes[a].count != nodes[b].count)
                        return nodes[a].count < nodes[b].count;
                    return nodes[a].seq < nodes[b].seq;
                });
                count_children(order);
                // how many nodes (from the back) get split before the list reaches N
                int total = prev, K = 0;
                for (int j = (int)order.size() - 1; j >= 0; j--) {
                    const QNode& q = nodes[order[j]];
                    int nonempty = (q.child[0] > 0) + (q.child[1] > 0) + (q.child[2] > 0) + (q.child[3] > 0);
                    total += nonempty - 1;
                    K++;
                    if (total >= N)
                        break;
                }
                std::vector<int> childOf2(nodes.size() * 4, -2);
                std::vector<int> front2;
                std::vector<char> erased(nodes.size(), 0);
                expandable.clear();
                for (int t = 0; t < K; t++) {
                    int id = order[order.size() - 1 - t];
                    int c[4];
                    split(id, c);
                    erased[id] = 1;
                    for (int k = 0; k < 4; k++) {
                        childOf2[id * 4 + k] = c[k];
                        if (c[k] >= 0) {
                            front2.push_back(c[k]);
                            if (nodes[c[k]].count > 1)
                                expandable.push_back(c[k]);
                        }
                    }
                }
                relabel(childOf2);
                std::vector<int> nl(front2.rbegin(), front2.rend());
                for (int id : list)
                    if (!erased[id])
                        nl.push_back(id);
                list.swap(nl);
                if ((int)list.size() >= N || (int)list.size() == prev)
                    finish = true;
            }
        }
    }

    // best point per node: max response, lowest candidate index on ties
    std::vector<int> best(nodes.size(), -1);
    for (int p = 0; p < n; p++) {
        int id = label[p];
        if (best[id] < 0 || cand_s(cand[p]) > cand_s(cand[best[id]]))
            best[id] = p;
    }
    out.reserve(list.size());
    for (int id : list)
        out.push_back(cand[best[id]]);
}

} // namespace orc

extern "C" int orc_quadtree_labels(const unsigned* cand, int n, int minX, int maxX, int minY, int maxY, int N, unsigned* out)
{
    std::vector<unsigned> o;
    orc::host_quadtree(cand, n, minX, maxX, minY, maxY, N, o);
    for (size_t i = 0; i < o.size(); i++)
        out[i] = o[i];
    return (int)o.size();
}
