// Box primitives (host side): six rectangles that are exactly the faces of one world-axis-aligned box — a Cube::new
// (src/geometry.rs:250-287) under a translation, the 400 ground blocks of examples/main.rs — are handed to the tree as ONE leaf item.
// A line crosses the surface of a convex box at most twice, so the closest hit among the six faces is the entry crossing, or the exit
// crossing when the entry face is the ray's own or lies behind the origin: one slab test (intersect.cuh box_hit) answers with the
// face's own rectangle (its primitive id, its plane, `t >= 1e-6`), and the tree holds one leaf instead of six.  The six rectangles stay
// device primitives (shade reads their records; a ray's originating primitive is the face), they are only no longer leaves.
//
// Detection is purely geometric and conservative: six rectangles with CONSECUTIVE DFS ranks (what Cube::new emits), each axis-aligned
// in the world frame, two per axis on distinct planes, every one covering its whole face of the common box (within a few f32 ulps).
// Anything else stays a rectangle.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

#include "../../include/rt_b200.h"

namespace rtb {

struct BoxGroup {
    uint32_t face[6];  // input primitive ids of the faces x-, x+, y-, y+, z-, z+
    float lo[3], hi[3];  // the faces' own planes
};

namespace boxdetail {
struct FaceRec {
    int axis;         // normal axis in the world frame, -1: not axis-aligned
    double w;         // plane: world coordinate `axis` of the rectangle
    double lo[3], hi[3];  // in-plane extent on the two other axes (entries of `axis` unused)
};
inline bool unitRow(const float* m, int r, int& k, double& sign) {
    k = -1;
    for (int c = 0; c < 3; ++c) {
        const double v = m[4 * r + c];
        if (std::fabs(std::fabs(v) - 1.0) <= 1e-6) {
            if (k >= 0) return false;
            k = c;
            sign = v > 0 ? 1.0 : -1.0;
        } else if (std::fabs(v) > 1e-6) return false;
    }
    return k >= 0;
}
// local = s * world_k + t per row of local_from_world: plane z = 0 <=> world_kz = -s_z t_z; |local x| <= hx <=> |world_kx - (-s_x t_x)| <= hx
inline FaceRec faceOf(const RtPrim& p) {
    FaceRec f{};
    f.axis = -1;
    if (p.kind != RT_PRIM_RECT) return f;
    int kx, ky, kz;
    double sx, sy, sz;
    const float* m = p.local_from_world;
    if (!unitRow(m, 0, kx, sx) || !unitRow(m, 1, ky, sy) || !unitRow(m, 2, kz, sz)) return f;
    if (kx == ky || kx == kz || ky == kz) return f;
    const double hx = 0.5 * std::fabs((double)p.param[0]), hy = 0.5 * std::fabs((double)p.param[1]);
    const double cx = -sx * (double)m[3], cy = -sy * (double)m[7];
    f.w = (double)(float)(-sz * (double)m[11]);
    f.lo[kx] = cx - hx; f.hi[kx] = cx + hx;
    f.lo[ky] = cy - hy; f.hi[ky] = cy + hy;
    f.lo[kz] = f.hi[kz] = f.w;
    f.axis = kz;
    return f;
}
inline bool near(double x, double y) { return std::fabs(x - y) <= 4e-7 * std::max(1.0, std::max(std::fabs(x), std::fabs(y))); }
}  // namespace boxdetail

// Groups of six input primitives that form a box; group_of[i] = index of the group of input primitive i, or -1.
inline std::vector<BoxGroup> find_box_groups(const RtPrim* prims, uint32_t n, std::vector<int>& group_of) {
    using namespace boxdetail;
    std::vector<BoxGroup> groups;
    group_of.assign(n, -1);
    std::vector<uint32_t> by_rank(n);
    for (uint32_t i = 0; i < n; ++i) by_rank[i] = i;
    std::stable_sort(by_rank.begin(), by_rank.end(), [&](uint32_t a, uint32_t b) { return prims[a].order < prims[b].order; });
    std::vector<FaceRec> rec(n);
    for (uint32_t i = 0; i < n; ++i) rec[i] = faceOf(prims[i]);
    for (uint32_t s = 0; s + 6 <= n;) {
        BoxGroup g{};
        int slot_of[6];
        bool ok = true;
        int seen[3] = {0, 0, 0};
        double plane[3][2];
        for (int j = 0; j < 6 && ok; ++j) {
            const uint32_t id = by_rank[s + j];
            const FaceRec& f = rec[id];
            ok = f.axis >= 0 && seen[f.axis] < 2 && prims[id].order == prims[by_rank[s]].order + (uint32_t)j;
            if (ok) plane[f.axis][seen[f.axis]++] = f.w;
        }
        double lo[3], hi[3];
        for (int k = 0; k < 3 && ok; ++k) {
            ok = seen[k] == 2 && !near(plane[k][0], plane[k][1]);
            lo[k] = std::min(plane[k][0], plane[k][1]);
            hi[k] = std::max(plane[k][0], plane[k][1]);
        }
        for (int j = 0; j < 6 && ok; ++j) {
            const FaceRec& f = rec[by_rank[s + j]];
            for (int a = 0; a < 3 && ok; ++a)
                if (a != f.axis) ok = near(f.lo[a], lo[a]) && near(f.hi[a], hi[a]);  // the rectangle IS the whole face
            slot_of[j] = 2 * f.axis + (f.w == hi[f.axis] ? 1 : 0);
        }
        if (!ok) { ++s; continue; }
        for (int j = 0; j < 6; ++j) g.face[slot_of[j]] = by_rank[s + j];
        for (int k = 0; k < 3; ++k) { g.lo[k] = (float)lo[k]; g.hi[k] = (float)hi[k]; }
        for (int j = 0; j < 6; ++j) group_of[by_rank[s + j]] = (int)groups.size();
        groups.push_back(g);
        s += 6;
    }
    return groups;
}

}  // namespace rtb
