// Binned-SAH BVH builder (host).  See bvh_build.h.
#include "bvh_build.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <atomic>
#include <numeric>
#include <thread>

namespace rtb {
namespace {

struct Grow {
    float lo[3], hi[3];
    Grow() { for (int a = 0; a < 3; ++a) { lo[a] = std::numeric_limits<float>::infinity(); hi[a] = -std::numeric_limits<float>::infinity(); } }
    void add(const Box3& b) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    void add(const Grow& b) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    void addPoint(const float* p) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); } }
    bool empty() const { return lo[0] > hi[0]; }
    double area() const {
        if (empty()) return 0.0;
        const double dx = (double)hi[0] - lo[0], dy = (double)hi[1] - lo[1], dz = (double)hi[2] - lo[2];
        return 2.0 * (dx * dy + dy * dz + dz * dx);
    }
};

double boxArea(const Box3& b) {
    const double dx = (double)b.hi[0] - b.lo[0], dy = (double)b.hi[1] - b.lo[1], dz = (double)b.hi[2] - b.lo[2];
    return 2.0 * (dx * dy + dy * dz + dz * dx);
}

struct TmpNode {
    Grow box;
    int32_t left = -1, right = -1;  // children (TmpNode indices) or -1
    uint32_t first = 0, count = 0;  // leaf range in the index array
};

struct Builder {
    const std::vector<Box3>& boxes;
    const BvhOptions& opt;
    std::vector<uint32_t> idx;       // permutation being partitioned (entries are input primitive ids)
    std::vector<float> cx, cy, cz;   // centroids by input id
    std::vector<TmpNode> tmp;
    uint32_t depth_seen = 0;
    // Subtrees of at most `defer_below` primitives are not built by the top-level recursion but recorded as tasks and built
    // concurrently afterwards (disjoint ranges of idx, thread-local node vectors spliced in task order: the tree is the same
    // for every thread count).  1 M spheres: 1.2 s -> a fraction of that on the box's 16 threads.
    struct Task { int32_t node; uint32_t first, count, depth; };
    std::vector<Task> tasks;
    uint32_t defer_below = 0;

    Builder(const std::vector<Box3>& b, const BvhOptions& o) : boxes(b), opt(o) {}

    const float* centroidAxis(int a) const { return a == 0 ? cx.data() : (a == 1 ? cy.data() : cz.data()); }

    int32_t build(std::vector<TmpNode>& tmp, uint32_t& depth_seen, uint32_t first, uint32_t count, uint32_t depth, bool may_defer) {
        depth_seen = std::max(depth_seen, depth);
        const int32_t me = (int32_t)tmp.size();
        tmp.emplace_back();
        if (may_defer && count <= defer_below && count > 1) {
            tasks.push_back(Task{me, first, count, depth});
            return me;
        }
        Grow box, cbox;
        for (uint32_t i = first; i < first + count; ++i) {
            const uint32_t p = idx[i];
            box.add(boxes[p]);
            const float c[3] = {cx[p], cy[p], cz[p]};
            cbox.addPoint(c);
        }
        tmp[me].box = box;
        auto makeLeaf = [&]() { tmp[me].first = first; tmp[me].count = count; return me; };
        if (count == 1) return makeLeaf();

        // binned SAH over the three axes
        const uint32_t B = opt.bins;
        double bestCost = std::numeric_limits<double>::infinity();
        int bestAxis = -1;
        uint32_t bestSplit = 0;
        std::vector<Grow> binBox(B), rightAcc(B);
        std::vector<uint32_t> binCnt(B);
        for (int a = 0; a < 3; ++a) {
            const float lo = cbox.lo[a], ext = cbox.hi[a] - cbox.lo[a];
            if (!(ext > 0.0f)) continue;
            const float scale = (float)B / ext;
            const float* ca = centroidAxis(a);
            for (uint32_t b = 0; b < B; ++b) { binBox[b] = Grow(); binCnt[b] = 0; }
            for (uint32_t i = first; i < first + count; ++i) {
                const uint32_t p = idx[i];
                uint32_t b = (uint32_t)std::max(0.0f, (ca[p] - lo) * scale);
                if (b >= B) b = B - 1;
                binBox[b].add(boxes[p]);
                binCnt[b]++;
            }
            Grow acc;
            for (uint32_t b = B; b-- > 1;) { acc.add(binBox[b]); rightAcc[b] = acc; }
            Grow left;
            uint32_t nl = 0;
            for (uint32_t b = 0; b + 1 < B; ++b) {
                left.add(binBox[b]);
                nl += binCnt[b];
                const uint32_t nr = count - nl;
                if (nl == 0 || nr == 0) continue;
                const double cost = left.area() * nl + rightAcc[b + 1].area() * nr;
                if (cost < bestCost) { bestCost = cost; bestAxis = a; bestSplit = b + 1; }
            }
        }
        const double parentArea = std::max(box.area(), 1e-30);
        const double splitCost = opt.traversal_cost + bestCost / parentArea;
        const bool forceSplit = count > opt.max_leaf;
        if (depth >= opt.max_depth) return makeLeaf();  // depth cap (keeps the traversal stack bounded); leaf may exceed max_leaf
        uint32_t mid;
        if (bestAxis < 0) {
            if (!forceSplit) return makeLeaf();
            mid = first + count / 2;  // identical centroids: split the range in half
        } else {
            if (!forceSplit && splitCost >= (double)count) return makeLeaf();
            const float lo = cbox.lo[bestAxis], ext = cbox.hi[bestAxis] - cbox.lo[bestAxis];
            const float scale = (float)B / ext;
            const float* ca = centroidAxis(bestAxis);
            auto it = std::partition(idx.begin() + first, idx.begin() + first + count, [&](uint32_t p) {
                uint32_t b = (uint32_t)std::max(0.0f, (ca[p] - lo) * scale);
                if (b >= B) b = B - 1;
                return b < bestSplit;
            });
            mid = (uint32_t)(it - idx.begin());
            if (mid == first || mid == first + count) mid = first + count / 2;
        }
        const int32_t l = build(tmp, depth_seen, first, mid - first, depth + 1, may_defer);
        const int32_t r = build(tmp, depth_seen, mid, first + count - mid, depth + 1, may_defer);
        tmp[me].left = l;
        tmp[me].right = r;
        return me;
    }
};

void emitNode(BvhNode32& out, const TmpNode& t, uint32_t ref, uint32_t count) {
    for (int a = 0; a < 3; ++a) { out.lo[a] = t.box.lo[a]; out.hi[a] = t.box.hi[a]; }
    out.left_or_first = ref;
    out.count = count;
}

}  // namespace

std::vector<uint32_t> select_always(const std::vector<Box3>& boxes, const BvhOptions& opt) {
    // primitives that span most of what is left (sky, ground, fog): always-tested prefix
    const uint32_t n = (uint32_t)boxes.size();
    std::vector<uint8_t> pulled(n, 0);
    std::vector<uint32_t> always;
    std::vector<uint32_t> byArea(n);
    std::iota(byArea.begin(), byArea.end(), 0u);
    const uint32_t keep = std::min<uint32_t>(n, opt.max_always + 1);
    std::partial_sort(byArea.begin(), byArea.begin() + keep, byArea.end(),
                      [&](uint32_t a, uint32_t b) { const double aa = boxArea(boxes[a]), ab = boxArea(boxes[b]); return aa > ab || (aa == ab && a < b); });
    for (uint32_t k = 0; k < keep && always.size() < opt.max_always; ++k) {
        const uint32_t cand = byArea[k];
        Grow rest;
        for (uint32_t i = 0; i < n; ++i)
            if (!pulled[i]) rest.add(boxes[i]);
        if (boxArea(boxes[cand]) >= opt.always_area_fraction * rest.area()) {
            pulled[cand] = 1;
            always.push_back(cand);
        } else {
            break;
        }
    }
    return always;
}

BvhResult build_bvh(const std::vector<Box3>& boxes, const BvhOptions& opt) {
    BvhResult res;
    const uint32_t n = (uint32_t)boxes.size();
    std::vector<uint8_t> pulled(n, 0);

    // 1. peel the always-tested prefix
    for (uint32_t id : select_always(boxes, opt)) {
        pulled[id] = 1;
        res.order.push_back(id);
        res.n_always++;
    }

    // 2. SAH tree over the rest
    Builder b(boxes, opt);
    b.cx.resize(n); b.cy.resize(n); b.cz.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
        b.cx[i] = 0.5f * (boxes[i].lo[0] + boxes[i].hi[0]);
        b.cy[i] = 0.5f * (boxes[i].lo[1] + boxes[i].hi[1]);
        b.cz[i] = 0.5f * (boxes[i].lo[2] + boxes[i].hi[2]);
        if (!pulled[i]) b.idx.push_back(i);
    }
    const uint32_t m = (uint32_t)b.idx.size();
    if (m == 0) return res;
    b.tmp.reserve(2 * (size_t)m);
    unsigned hw = std::thread::hardware_concurrency();
    unsigned n_threads = std::max(1u, std::min(hw ? hw : 1u, 32u));
    if (opt.threads >= 1) n_threads = std::min(opt.threads, 64u);
    b.defer_below = (n_threads > 1 && m >= 65536u) ? m / (8u * n_threads) : 0u;
    const int32_t root = b.build(b.tmp, b.depth_seen, 0, m, 0, b.defer_below > 0);
    if (!b.tasks.empty()) {
        std::vector<std::vector<TmpNode>> local(b.tasks.size());
        std::vector<uint32_t> local_depth(b.tasks.size(), 0u);
        std::atomic<size_t> next{0};
        auto worker = [&]() {
            for (size_t t = next.fetch_add(1); t < b.tasks.size(); t = next.fetch_add(1)) {
                const Builder::Task& k = b.tasks[t];
                local[t].reserve(2 * (size_t)k.count);
                b.build(local[t], local_depth[t], k.first, k.count, k.depth, false);
            }
        };
        std::vector<std::thread> pool;
        for (unsigned i = 1; i < n_threads; ++i) pool.emplace_back(worker);
        worker();
        for (auto& th : pool) th.join();
        for (size_t t = 0; t < b.tasks.size(); ++t) {  // splice: local node 0 replaces the placeholder, local k >= 1 lands at off + k
            const int32_t off = (int32_t)b.tmp.size() - 1;
            std::vector<TmpNode>& loc = local[t];
            for (TmpNode& nd : loc) { if (nd.left >= 0) { nd.left += off; nd.right += off; } }
            b.tmp[b.tasks[t].node] = loc[0];
            b.tmp.insert(b.tmp.end(), loc.begin() + 1, loc.end());
            b.depth_seen = std::max(b.depth_seen, local_depth[t]);
        }
    }
    res.depth = b.depth_seen;

    // 3. breadth-first emission with adjacent sibling pairs; leaf ranges are offset by the always-list
    for (uint32_t i = 0; i < m; ++i) res.order.push_back(b.idx[i]);
    const uint32_t base = res.n_always;
    res.nodes.resize(2);
    std::memset(res.nodes.data(), 0, 2 * sizeof(BvhNode32));
    const TmpNode& rt = b.tmp[root];
    if (rt.left < 0) {
        res.root_is_leaf = true;
        emitNode(res.nodes[0], rt, base + rt.first, rt.count);
        return res;
    }
    // Emission order.  The first treelet is the top of the tree, breadth-first (extend stages that prefix in shared memory);
    // below it the tree is cut into treelets of `treelet_pairs` sibling pairs (2 KB), each stored contiguously and visited
    // depth-first, so that the node fetches of one ray's descent fall into a few cache lines instead of one line per level
    // (plain breadth-first order: 65 % L1 / 75 % L2 hit rate on the 1 M-sphere tree, r1j).
    const uint32_t top_pairs = std::max<uint32_t>(1u, opt.top_nodes / 2u), treelet_pairs = std::max<uint32_t>(1u, opt.treelet_nodes / 2u);
    std::vector<std::pair<int32_t, uint32_t>> roots;  // treelet roots still to emit (tmp node, emitted slot), LIFO
    roots.emplace_back(root, 0u);
    std::vector<std::pair<int32_t, uint32_t>> queue;
    bool first_treelet = true;
    while (!roots.empty()) {
        queue.clear();
        queue.push_back(roots.back());
        roots.pop_back();
        const uint32_t budget = first_treelet ? top_pairs : treelet_pairs;
        first_treelet = false;
        uint32_t used = 0;
        size_t head = 0;
        for (; head < queue.size() && used < budget; ++head, ++used) {
            const int32_t t = queue[head].first;
            const uint32_t slot = queue[head].second;
            const TmpNode& node = b.tmp[t];
            const uint32_t pairAt = (uint32_t)res.nodes.size();
            res.nodes.resize(res.nodes.size() + 2);
            emitNode(res.nodes[slot], node, pairAt, 0u);
            const int32_t kids[2] = {node.left, node.right};
            for (int c = 0; c < 2; ++c) {
                const TmpNode& kid = b.tmp[kids[c]];
                if (kid.left < 0) emitNode(res.nodes[pairAt + c], kid, base + kid.first, kid.count);
                else queue.emplace_back(kids[c], pairAt + c);
            }
        }
        for (size_t i = queue.size(); i-- > head;) roots.push_back(queue[i]);  // the frontier becomes new treelet roots (left subtree first)
    }
    return res;
}

}  // namespace rtb
