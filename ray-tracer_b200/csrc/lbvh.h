// Device-side LBVH build (lbvh.cu): Morton sort + Karras radix tree + bottom-up fit, emitted in the sibling-pair layout of bvh_build.h.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <vector>

#include "bvh_build.h"

namespace rtb {

struct LbvhResult {
    float4* d_nodes = nullptr;    // device array of 2 * n_nodes float4, cudaMalloc'ed: the caller owns it
    uint32_t n_nodes = 0;         // node 0 = root, node 1 = padding, sibling pairs at even indices
    uint32_t depth = 0;
    std::vector<uint32_t> order;  // sorted position -> index into `boxes`
};

// `boxes`: world-space boxes of the primitives that go into the tree (the always-list already removed); leaf ranges are
// emitted as leaf_base + sorted position.  Needs boxes.size() > max_leaf.
cudaError_t build_lbvh_device(const std::vector<Box3>& boxes, uint32_t leaf_base, uint32_t max_leaf, cudaStream_t st, LbvhResult& out);

}  // namespace rtb
