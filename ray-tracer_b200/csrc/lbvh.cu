// Device-side BVH build: a linear BVH over the flattened primitives, built entirely on the GPU and emitted
// in the same 32-byte sibling-pair node layout the extend kernels traverse (bvh_build.h).  It replaces, as an option, the host
// binned-SAH rebuild of rt_scene_create — the counterpart of the reference's BoundingVolumeHierarchyNode::new
// (src/optimize.rs:363-461: random axis, median split, `bound()` recomputed inside the sort comparator).
//
//   1. k_morton       30-bit Morton code of every box centroid inside the scene bounds
//   2. cub radix sort  (keys = codes, values = primitive ids; stable, so equal codes keep id order)
//   3. k_radix_tree   Karras 2012, "Maximizing Parallelism in the Construction of BVHs, Octrees, and k-d Trees": every internal
//                     node finds its key range and split in parallel; equal codes fall back to the index bits
//   4. k_fit          bottom-up boxes, one thread per leaf, the second arrival at a node (atomic flag) merges the children
//   5. cub scan + k_emit  internal nodes that cover more than `max_leaf` primitives become interior nodes (rank -> pair slot);
//                     anything smaller is emitted as ONE leaf over its contiguous range of the sorted order
// Closest-hit answers do not depend on topology (tests compare against the SAH tree); the tree is of lower quality than the SAH
// one (more box tests per ray) and its top levels are not a prefix of the node array, so extend stages only the root pair.
#include <cuda_runtime.h>

#include <cub/cub.cuh>

#include <cstdint>
#include <vector>

#include "bvh_build.h"
#include "lbvh.h"

namespace rtb {
namespace {

__device__ __forceinline__ uint32_t expand_bits(uint32_t v) {  // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void k_morton(const Box3* __restrict__ boxes, uint32_t m, float3 lo, float3 inv_ext, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const Box3 b = boxes[i];
    const float cx = (0.5f * (b.lo[0] + b.hi[0]) - lo.x) * inv_ext.x, cy = (0.5f * (b.lo[1] + b.hi[1]) - lo.y) * inv_ext.y,
                cz = (0.5f * (b.lo[2] + b.hi[2]) - lo.z) * inv_ext.z;
    const uint32_t x = (uint32_t)fminf(fmaxf(cx * 1024.0f, 0.0f), 1023.0f), y = (uint32_t)fminf(fmaxf(cy * 1024.0f, 0.0f), 1023.0f),
                   z = (uint32_t)fminf(fmaxf(cz * 1024.0f, 0.0f), 1023.0f);
    keys[i] = (expand_bits(x) << 2) | (expand_bits(y) << 1) | expand_bits(z);
    vals[i] = i;
}

// length of the common prefix of keys i and j (index bits break ties), -1 outside the array
__device__ __forceinline__ int delta(const uint32_t* __restrict__ keys, int m, int i, int j) {
    if (j < 0 || j >= m) return -1;
    const uint32_t a = keys[i], b = keys[j];
    if (a == b) return 32 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clz(a ^ b);
}

// Node ids: internal 0 .. m-2, leaf k = m-1+k.  range[i] = (first, last) of internal node i in the sorted order.
__global__ void k_radix_tree(const uint32_t* __restrict__ keys, int m, int2* __restrict__ children, int2* __restrict__ range, int* __restrict__ parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m - 1) return;
    const int d = (delta(keys, m, i, i + 1) - delta(keys, m, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(keys, m, i, i - d);
    int lmax = 2;
    while (delta(keys, m, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta(keys, m, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, m, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) / 2;
        if (delta(keys, m, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    const int left = (first == gamma) ? (m - 1 + gamma) : gamma;
    const int right = (last == gamma + 1) ? (m - 1 + gamma + 1) : (gamma + 1);
    children[i] = make_int2(left, right);
    range[i] = make_int2(first, last);
    parent[left] = i;
    parent[right] = i;
    if (i == 0) parent[0] = -1;
}

struct NodeBox {
    float lo[3], hi[3];
};

__global__ void k_fit(const Box3* __restrict__ boxes, const uint32_t* __restrict__ vals, int m, const int2* __restrict__ children, const int* __restrict__ parent,
                      NodeBox* node_box, int* flags, int* max_depth) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m) return;
    const Box3 b = boxes[vals[k]];
    NodeBox nb;
    for (int a = 0; a < 3; ++a) { nb.lo[a] = b.lo[a]; nb.hi[a] = b.hi[a]; }
    int node = m - 1 + k, depth = 0;
    node_box[node] = nb;
    __threadfence();
    int cur = parent[node];
    while (cur >= 0) {
        ++depth;
        if (atomicAdd(&flags[cur], 1) == 0) break;  // first arrival: the sibling subtree is not finished yet
        const int2 ch = children[cur];
        const volatile NodeBox* l = node_box + ch.x;
        const volatile NodeBox* r = node_box + ch.y;
        for (int a = 0; a < 3; ++a) { nb.lo[a] = fminf(l->lo[a], r->lo[a]); nb.hi[a] = fmaxf(l->hi[a], r->hi[a]); }
        node_box[cur] = nb;
        __threadfence();
        cur = parent[cur];
    }
    // depth of this leaf (walk on to the root without touching flags)
    while (cur >= 0) { cur = parent[cur]; if (cur >= 0) ++depth; }
    atomicMax(max_depth, depth);
}

__global__ void k_interior_flags(const int2* __restrict__ range, int m, int max_leaf, uint32_t* __restrict__ flag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m - 1) return;
    const int2 r = range[i];
    flag[i] = (r.y - r.x + 1 > max_leaf) ? 1u : 0u;
}

// nodes: float4 pairs {lo.xyz, leftOrFirst} {hi.xyz, count}; node 0 = root, node 1 = padding, interior node of rank r owns the
// sibling pair at 2 (r + 1).
__global__ void k_emit(const int2* __restrict__ children, const int2* __restrict__ range, const NodeBox* __restrict__ node_box, const uint32_t* __restrict__ flag,
                       const uint32_t* __restrict__ rank, int m, uint32_t leaf_base, float4* __restrict__ nodes) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m - 1 || !flag[i]) return;
    const uint32_t pair_at = 2u * (rank[i] + 1u);
    if (i == 0) {
        const NodeBox b = node_box[0];
        nodes[0] = make_float4(b.lo[0], b.lo[1], b.lo[2], __uint_as_float(pair_at));
        nodes[1] = make_float4(b.hi[0], b.hi[1], b.hi[2], __uint_as_float(0u));
        nodes[2] = nodes[3] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const int2 ch = children[i];
    const int kids[2] = {ch.x, ch.y};
    for (int c = 0; c < 2; ++c) {
        const int kid = kids[c];
        const NodeBox b = node_box[kid];
        uint32_t ref, count;
        if (kid < m - 1 && flag[kid]) { ref = 2u * (rank[kid] + 1u); count = 0u; }
        else if (kid < m - 1) { const int2 r = range[kid]; ref = leaf_base + (uint32_t)r.x; count = (uint32_t)(r.y - r.x + 1); }
        else { ref = leaf_base + (uint32_t)(kid - (m - 1)); count = 1u; }
        nodes[2u * (pair_at + c)] = make_float4(b.lo[0], b.lo[1], b.lo[2], __uint_as_float(ref));
        nodes[2u * (pair_at + c) + 1u] = make_float4(b.hi[0], b.hi[1], b.hi[2], __uint_as_float(count));
    }
}

template <class T>
struct Dev {
    T* p = nullptr;
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, (n ? n : 1) * sizeof(T)); }
    ~Dev() { if (p) cudaFree(p); }
};

}  // namespace

cudaError_t build_lbvh_device(const std::vector<Box3>& boxes, uint32_t leaf_base, uint32_t max_leaf, cudaStream_t st, LbvhResult& out) {
    const uint32_t m = (uint32_t)boxes.size();
    out = LbvhResult();
    if (m < 2 || m <= max_leaf || m > (1u << 30)) return cudaErrorInvalidValue;
    // scene bounds of the centroids (host: one pass over data that is being uploaded anyway)
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (const Box3& b : boxes)
        for (int a = 0; a < 3; ++a) {
            const float c = 0.5f * (b.lo[a] + b.hi[a]);
            lo[a] = fminf(lo[a], c);
            hi[a] = fmaxf(hi[a], c);
        }
    float3 inv_ext;
    inv_ext.x = hi[0] > lo[0] ? 1.0f / (hi[0] - lo[0]) : 0.0f;
    inv_ext.y = hi[1] > lo[1] ? 1.0f / (hi[1] - lo[1]) : 0.0f;
    inv_ext.z = hi[2] > lo[2] ? 1.0f / (hi[2] - lo[2]) : 0.0f;

    Dev<Box3> d_boxes;
    Dev<uint32_t> keys_a, keys_b, vals_a, vals_b, flag, rank;
    Dev<int2> children, range;
    Dev<int> parent, flags, depth;
    Dev<NodeBox> node_box;
    Dev<uint8_t> temp;
    cudaError_t e;
#define LB(call) do { e = (call); if (e != cudaSuccess) return e; } while (0)
    LB(d_boxes.alloc(m)); LB(keys_a.alloc(m)); LB(keys_b.alloc(m)); LB(vals_a.alloc(m)); LB(vals_b.alloc(m));
    LB(flag.alloc(m)); LB(rank.alloc(m)); LB(children.alloc(m)); LB(range.alloc(m)); LB(parent.alloc(2 * (size_t)m)); LB(flags.alloc(m));
    LB(depth.alloc(1)); LB(node_box.alloc(2 * (size_t)m));
    LB(cudaMemcpyAsync(d_boxes.p, boxes.data(), (size_t)m * sizeof(Box3), cudaMemcpyHostToDevice, st));
    LB(cudaMemsetAsync(flags.p, 0, (size_t)m * sizeof(int), st));
    LB(cudaMemsetAsync(depth.p, 0, sizeof(int), st));
    const uint32_t T = 256, G = (m + T - 1) / T;
    k_morton<<<G, T, 0, st>>>(d_boxes.p, m, make_float3(lo[0], lo[1], lo[2]), inv_ext, keys_a.p, vals_a.p);
    size_t sort_bytes = 0, scan_bytes = 0;
    LB(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, keys_a.p, keys_b.p, vals_a.p, vals_b.p, (int)m, 0, 30, st));
    LB(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, flag.p, rank.p, (int)m, st));
    LB(temp.alloc(sort_bytes > scan_bytes ? sort_bytes : scan_bytes));
    LB(cub::DeviceRadixSort::SortPairs(temp.p, sort_bytes, keys_a.p, keys_b.p, vals_a.p, vals_b.p, (int)m, 0, 30, st));
    k_radix_tree<<<G, T, 0, st>>>(keys_b.p, (int)m, children.p, range.p, parent.p);
    k_fit<<<G, T, 0, st>>>(d_boxes.p, vals_b.p, (int)m, children.p, parent.p, node_box.p, flags.p, depth.p);
    k_interior_flags<<<G, T, 0, st>>>(range.p, (int)m, (int)max_leaf, flag.p);
    LB(cudaMemsetAsync(flag.p + (m - 1), 0, sizeof(uint32_t), st));  // entry m-1 is not an internal node
    LB(cub::DeviceScan::ExclusiveSum(temp.p, scan_bytes, flag.p, rank.p, (int)m, st));
    uint32_t n_interior = 0;
    int h_depth = 0;
    LB(cudaMemcpyAsync(&n_interior, rank.p + (m - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, st));  // exclusive sum at m-1 = total
    LB(cudaMemcpyAsync(&h_depth, depth.p, sizeof(int), cudaMemcpyDeviceToHost, st));
    LB(cudaStreamSynchronize(st));
    out.n_nodes = 2u * (n_interior + 1u);
    out.depth = (uint32_t)h_depth;
    float4* nodes = nullptr;
    LB(cudaMalloc(&nodes, (size_t)out.n_nodes * 2 * sizeof(float4)));
    k_emit<<<G, T, 0, st>>>(children.p, range.p, node_box.p, flag.p, rank.p, (int)m, leaf_base, nodes);
    out.order.resize(m);
    e = cudaMemcpyAsync(out.order.data(), vals_b.p, (size_t)m * sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) { cudaFree(nodes); return e; }
#undef LB
    out.d_nodes = nodes;
    return cudaSuccess;
}

}  // namespace rtb
