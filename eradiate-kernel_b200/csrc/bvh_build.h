/*
 * bvh_build.h -- host-side binned-SAH BVH over the scene-wide triangle pool.
 *
 * Replaces the reference's SAH kd-tree build (include/mitsuba/render/kdtree.h,
 * src/librender/scene_native.inl:3-10) for the device path.  Only the traversal
 * RESULT has to agree with the reference (closest hit); the tree
 * itself is free.  Boxes are padded so that the float slab test used on the
 * device is conservative with respect to the exact triangle test.
 */
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

#include "vp_types.cuh"

namespace VP_NS {

struct BuildTri { float bmin[3], bmax[3], c[3]; uint32_t shape, face; };

/* The device traversal (vp_scene.cuh, ray_intersect) pushes at most one node per tree level. */
constexpr uint32_t BVH_STACK_CAPACITY = 64;
constexpr uint32_t BVH_FORCE_MEDIAN_DEPTH = 24;

struct BvhStats {
    uint32_t n_nodes = 0, n_leaves = 0;
    uint32_t max_depth = 0;              /* depth of the deepest leaf (root = 0) = largest possible stack use */
    uint32_t n_forced_median = 0;        /* inner nodes split by the index median because depth >= 24 */
    uint32_t n_degenerate_median = 0;    /* inner nodes split by the index median because SAH found no split */
};

inline void bvh_build(std::vector<BuildTri> &tris, float pad, std::vector<BvhNode> &nodes,
                      std::vector<BvhPrim> &prims, int max_leaf = 4, BvhStats *stats = nullptr) {
    nodes.clear();
    prims.clear();
    BvhStats st;
    if (stats) *stats = st;
    if (tris.empty()) return;
    const int NB = 16, MAX_LEAF = std::min(std::max(max_leaf, 1), 7);
    nodes.reserve(tris.size());
    prims.reserve(tris.size());
    struct Task { uint32_t node, lo, hi, depth; };
    std::vector<Task> stack;
    nodes.push_back(BvhNode{});
    stack.push_back({ 0u, 0u, (uint32_t) tris.size(), 0u });
    while (!stack.empty()) {
        Task t = stack.back();
        stack.pop_back();
        float bmin[3] = { INFINITY, INFINITY, INFINITY }, bmax[3] = { -INFINITY, -INFINITY, -INFINITY };
        float cmin[3] = { INFINITY, INFINITY, INFINITY }, cmax[3] = { -INFINITY, -INFINITY, -INFINITY };
        for (uint32_t i = t.lo; i < t.hi; ++i)
            for (int a = 0; a < 3; ++a) {
                bmin[a] = std::min(bmin[a], tris[i].bmin[a]); bmax[a] = std::max(bmax[a], tris[i].bmax[a]);
                cmin[a] = std::min(cmin[a], tris[i].c[a]);    cmax[a] = std::max(cmax[a], tris[i].c[a]);
            }
        BvhNode &n = nodes[t.node];
        for (int a = 0; a < 3; ++a) { n.bmin[a] = bmin[a] - pad; n.bmax[a] = bmax[a] + pad; }
        uint32_t cnt = t.hi - t.lo;
        auto make_leaf = [&]() {
            st.n_leaves++;
            st.max_depth = std::max(st.max_depth, t.depth);
            nodes[t.node].child = (uint32_t) prims.size();
            nodes[t.node].count = cnt;
            for (uint32_t i = t.lo; i < t.hi; ++i) prims.push_back(BvhPrim{ tris[i].shape, tris[i].face });
        };
        /* Leaves hold at most MAX_LEAF triangles (the device packs the count into 3 bits).  The device traversal
           stack holds 64 entries, one per level: SAH splits may be lopsided, so from depth 24 on the split is the
           index median, which bounds the depth by 24 + log2(n). */
        if (cnt <= (uint32_t) MAX_LEAF) { make_leaf(); continue; }
        const bool force_median = t.depth >= BVH_FORCE_MEDIAN_DEPTH;

        /* binned SAH over the three axes */
        int best_axis = -1, best_split = -1;
        float best_cost = INFINITY;
        for (int a = 0; a < 3 && !force_median; ++a) {
            float ext = cmax[a] - cmin[a];
            if (!(ext > 0.f)) continue;
            float scale = NB / ext;
            float bbmin[NB][3], bbmax[NB][3];
            uint32_t bcnt[NB];
            for (int b = 0; b < NB; ++b) {
                bcnt[b] = 0;
                for (int k = 0; k < 3; ++k) { bbmin[b][k] = INFINITY; bbmax[b][k] = -INFINITY; }
            }
            for (uint32_t i = t.lo; i < t.hi; ++i) {
                int b = std::min(NB - 1, (int) ((tris[i].c[a] - cmin[a]) * scale));
                bcnt[b]++;
                for (int k = 0; k < 3; ++k) {
                    bbmin[b][k] = std::min(bbmin[b][k], tris[i].bmin[k]);
                    bbmax[b][k] = std::max(bbmax[b][k], tris[i].bmax[k]);
                }
            }
            auto area = [](const float *mn, const float *mx) {
                float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
                return dx * dy + dy * dz + dz * dx;
            };
            float right_area[NB];
            uint32_t right_cnt[NB];
            float mn[3] = { INFINITY, INFINITY, INFINITY }, mx[3] = { -INFINITY, -INFINITY, -INFINITY };
            uint32_t c = 0;
            for (int b = NB - 1; b > 0; --b) {
                c += bcnt[b];
                for (int k = 0; k < 3; ++k) { mn[k] = std::min(mn[k], bbmin[b][k]); mx[k] = std::max(mx[k], bbmax[b][k]); }
                right_area[b] = c ? area(mn, mx) : 0.f;
                right_cnt[b] = c;
            }
            for (int k = 0; k < 3; ++k) { mn[k] = INFINITY; mx[k] = -INFINITY; }
            c = 0;
            for (int b = 0; b < NB - 1; ++b) {
                c += bcnt[b];
                for (int k = 0; k < 3; ++k) { mn[k] = std::min(mn[k], bbmin[b][k]); mx[k] = std::max(mx[k], bbmax[b][k]); }
                if (c == 0 || right_cnt[b + 1] == 0) continue;
                float cost = area(mn, mx) * c + right_area[b + 1] * right_cnt[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_split = b; }
            }
        }
        uint32_t mid;
        if (best_axis < 0) {
            /* all centroids coincide, or the tree is getting deep: split by index */
            mid = t.lo + cnt / 2;
            if (force_median) st.n_forced_median++; else st.n_degenerate_median++;
        } else {
            float ext = cmax[best_axis] - cmin[best_axis], scale = NB / ext, c0 = cmin[best_axis];
            int ax = best_axis, sp = best_split;
            auto it = std::partition(tris.begin() + t.lo, tris.begin() + t.hi, [&](const BuildTri &q) {
                int b = std::min(NB - 1, (int) ((q.c[ax] - c0) * scale));
                return b <= sp;
            });
            mid = (uint32_t) (it - tris.begin());
            if (mid == t.lo || mid == t.hi) mid = t.lo + cnt / 2;
        }
        uint32_t child = (uint32_t) nodes.size();
        nodes.push_back(BvhNode{});
        nodes.push_back(BvhNode{});
        nodes[t.node].child = child;
        nodes[t.node].count = 0;
        stack.push_back({ child + 1, mid, t.hi, t.depth + 1 });
        stack.push_back({ child, t.lo, mid, t.depth + 1 });
    }
    st.n_nodes = (uint32_t) nodes.size();
    if (stats) *stats = st;
}

} // namespace VP_NS
