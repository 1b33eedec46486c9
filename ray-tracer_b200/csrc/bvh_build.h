// Host-side builder of the device BVH: binned-SAH tree over the world-space boxes of the flattened
// primitives, emitted (top breadth-first, then in depth-first treelets) as 32-byte {lo, leftOrFirst}{hi, count} nodes with sibling pairs
// adjacent.  The reference builds a random-axis median-split tree (src/optimize.rs:363-440) whose
// closest-hit answers do not depend on topology, so the device tree is rebuilt for quality
// (SURVEY §7.2-4); primitives whose box spans most of the scene (sky / ground / fog spheres) are
// pulled out into an always-tested prefix instead of poisoning the top of the tree.
#pragma once
#include <cstdint>
#include <vector>

namespace rtb {

struct Box3 {
    float lo[3], hi[3];
};

struct BvhNode32 {  // 32 B, uploaded verbatim as two float4
    float lo[3];
    uint32_t left_or_first;
    float hi[3];
    uint32_t count;  // 0 = interior (children at left_or_first, left_or_first + 1)
};
static_assert(sizeof(BvhNode32) == 32, "node must be 32 bytes");

struct BvhResult {
    std::vector<BvhNode32> nodes;   // nodes[0] = root, nodes[1] = padding, pairs at even indices
    std::vector<uint32_t> order;    // device primitive index -> input primitive index (always-list first)
    uint32_t n_always = 0;
    uint32_t depth = 0;
    bool root_is_leaf = false;
};

struct BvhOptions {
    uint32_t max_leaf = 4;
    uint32_t bins = 16;
    float traversal_cost = 1.2f;   // cost of a box-pair test relative to one primitive test
    float always_area_fraction = 0.5f;
    uint32_t max_always = 8;
    uint32_t max_depth = 56;
    uint32_t top_nodes = 256;      // breadth-first prefix (what extend stages in shared memory for large trees)
    uint32_t treelet_nodes = 64;   // below the prefix: nodes per contiguously stored treelet (2 KB)
    uint32_t threads = 0;          // worker threads of the build (0 = one per host core, at most 32); the tree does not depend on it
};

BvhResult build_bvh(const std::vector<Box3>& boxes, const BvhOptions& opt = BvhOptions());
// Input ids of the primitives that go to the always-tested list (largest first); shared with the device builder (lbvh.cu).
std::vector<uint32_t> select_always(const std::vector<Box3>& boxes, const BvhOptions& opt);

}  // namespace rtb
