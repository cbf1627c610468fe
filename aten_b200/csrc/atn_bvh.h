// atn_bvh.h -- host-side re-encoding of the reference's threaded node lists into the traversal
// layout the sm_100a kernels walk (DESIGN.md "Data layout in HBM").
//
// Input  : std::vector<std::vector<GPUBvhNode>> exactly as the reference's builders emit them
//          (48-byte nodes, hit/miss links stored as floats; list 0 = top layer over instances,
//          list k = SBVH of object k, one triangle per leaf; reference
//          src/libaten/accelerator/threaded_bvh.h:13-56, sbvh.h:19-66, sbvh.cpp:827-950,
//          threaded_bvh.cpp:254-357).
// Output : * PairNode[]  64 B, 64-byte aligned: both children's boxes + both child references, so one
//            node fetch (4 x LDG.128, two 32 B sectors) decides both children;
//          * TriLeaf[]   48 B: v0 and the two edges v1-v0, v2-v0 gathered next to each other with
//            primid / mtrlid / mesh_id in the .w lanes (no index -> vertex indirection);
//          * InstLeaf[]  16 B: objid, root reference of the bottom tree, W2L matrix index, meshid.
//          Leaves are numbered in the reference's depth-first (hit-link) order; the index is the
//          tie-break rank that reproduces "first encountered wins" (SURVEY Q2).
// Boxes stay fp32 and un-quantised: the reference's slab test (aabb.h:61-86, +1e-6 bias) is part of
// its result, and primitive ids must match bit for bit (see DESIGN.md for why quantisation was
// rejected for this round).
#pragma once

#include <stdint.h>
#include <string.h>
#include <algorithm>
#include <cmath>
#include <string>
#include <vector>

#include "../../include/aten_b200.h"

namespace atn {

struct alignas(64) PairNode {
    // n0 = L.min.x, L.max.x, L.min.y, L.max.y ; n1 = R.min.x, R.max.x, R.min.y, R.max.y
    // n2 = L.min.z, L.max.z, R.min.z, R.max.z ; n3 = left ref, right ref, 0, 0 (as int bits)
    float n0[4], n1[4], n2[4];
    int32_t left, right, pad0, pad1;
};
static_assert(sizeof(PairNode) == 64, "PairNode");

struct alignas(16) TriLeaf {
    float v0[3]; int32_t primid;
    float e1[3]; int32_t mtrlid;      // v1 - v0 (the float subtraction intersectTriangle starts with)
    float e2[3]; int32_t mesh_id;     // v2 - v0
};
static_assert(sizeof(TriLeaf) == 48, "TriLeaf");

struct alignas(16) InstLeaf {
    int32_t objid;      // object index recorded in isect.objid
    int32_t root;       // reference of the bottom tree's root (see kRef*)
    int32_t w2l;        // index of the world-to-local matrix, -1: none (ray used as is)
    int32_t meshid;
};

// Child / root references: >= 0 inner PairNode index; < 0 leaf number ~ref; two reserved values.
constexpr int32_t kRefNull = INT32_MIN + 1;   // leaf that holds nothing traversable (e.g. a sphere)
constexpr int32_t kRefExit = INT32_MIN;       // stack marker: leave the current instance

struct EncodedBvh {
    std::vector<PairNode> nodes;
    std::vector<TriLeaf> tris;
    std::vector<InstLeaf> insts;
    int32_t top_root{ kRefNull };
    int32_t max_depth{ 0 };
    std::string error;
};

inline int32_t leaf_ref(int32_t leaf_index) { return ~leaf_index; }

namespace detail {

struct ListBuilder {
    const AtnBvhNode* list;
    uint32_t count;
    bool top;
    const AtnSceneBlob* scene;
    EncodedBvh* out;
    const std::vector<int32_t>* list_roots;   // for the top layer: root refs of the bottom lists
    int32_t depth_seen{ 0 };

    static bool is_leaf(const AtnBvhNode& n) { return n.object_id >= 0 || n.primid >= 0; }

    int32_t make_leaf(const AtnBvhNode& n)
    {
        if (top) {
            if (!(n.exid >= 0)) return kRefNull;        // no external tree: sphere / cube, never hit (F5)
            uint32_t bits; memcpy(&bits, &n.exid, 4);
            const int32_t exid = (int32_t)(bits & 0x7fffu);   // mainExid:15 (threaded_bvh.h:29-37)
            if (exid <= 0 || exid >= (int32_t)list_roots->size()) return kRefNull;
            InstLeaf il;
            il.objid = (int32_t)n.object_id;
            il.root = (*list_roots)[exid];
            const AtnObjectParameter& obj = scene->objects[(uint32_t)n.object_id];
            il.w2l = obj.mtx_id >= 0 ? obj.mtx_id + 1 : -1;    // threaded_bvh_traverser.h:149-158
            il.meshid = (int32_t)n.meshid;
            out->insts.push_back(il);
            return leaf_ref((int32_t)out->insts.size() - 1);
        }
        if (!(n.primid >= 0)) return kRefNull;
        const uint32_t primid = (uint32_t)n.primid;
        if (primid >= scene->triangle_count) { out->error = "triangle index out of range in node list"; return kRefNull; }
        const AtnTriangleParameter& tp = scene->triangles[primid];
        TriLeaf tl;
        for (int k = 0; k < 3; k++) {
            if ((uint32_t)tp.idx[k] >= scene->vertex_count) { out->error = "vertex index out of range"; return kRefNull; }
        }
        const AtnVec4& a = scene->positions[tp.idx[0]];
        const AtnVec4& b = scene->positions[tp.idx[1]];
        const AtnVec4& c = scene->positions[tp.idx[2]];
        tl.v0[0] = a.x; tl.v0[1] = a.y; tl.v0[2] = a.z; tl.primid = (int32_t)primid;
        tl.e1[0] = b.x - a.x; tl.e1[1] = b.y - a.y; tl.e1[2] = b.z - a.z; tl.mtrlid = tp.mtrlid;
        tl.e2[0] = c.x - a.x; tl.e2[1] = c.y - a.y; tl.e2[2] = c.z - a.z; tl.mesh_id = tp.mesh_id;
        out->tris.push_back(tl);
        return leaf_ref((int32_t)out->tris.size() - 1);
    }

    // Conservative bounds of a leaf: the triangle's own AABB, or the AABB of the instance's root box
    // (bottom list node 0) carried to world space by its L2W matrix.
    void leaf_bounds(int32_t ref, float mn[3], float mx[3]) const
    {
        const int32_t idx = ~ref;
        if (!top) {
            const TriLeaf& t = out->tris[idx];
            for (int a = 0; a < 3; a++) {
                const float p0 = t.v0[a], p1 = t.v0[a] + t.e1[a], p2 = t.v0[a] + t.e2[a];
                // v0 + (v1 - v0) may round away from v1 by an ulp: widen by one ulp-ish step
                const float lo = std::min(p0, std::min(p1, p2)), hi = std::max(p0, std::max(p1, p2));
                const float pad = 4e-7f * std::max(std::fabs(lo), std::fabs(hi)) + 1e-30f;
                mn[a] = lo - pad; mx[a] = hi + pad;
            }
            return;
        }
        const InstLeaf& il = out->insts[idx];
        const AtnObjectParameter& obj = scene->objects[il.objid];
        // find the bottom list of this instance again: its root box is node 0 of that list
        const AtnBvhNode* root = nullptr;
        for (uint32_t k = 1; k < scene->node_list_count; k++) {
            if ((*list_roots)[k] == il.root && scene->node_list_sizes[k] > 0) { root = &scene->node_lists[k][0]; break; }
        }
        const float big = 3.0e38f;
        if (!root) { for (int a = 0; a < 3; a++) { mn[a] = -big; mx[a] = big; } return; }
        float M[16] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
        if (obj.mtx_id >= 0 && (uint32_t)obj.mtx_id < scene->matrix_count) memcpy(M, scene->matrices[obj.mtx_id].m, sizeof(M));
        for (int a = 0; a < 3; a++) { mn[a] = big; mx[a] = -big; }
        for (int c = 0; c < 8; c++) {
            const double x = (c & 1) ? root->boxmax[0] : root->boxmin[0];
            const double y = (c & 2) ? root->boxmax[1] : root->boxmin[1];
            const double z = (c & 4) ? root->boxmax[2] : root->boxmin[2];
            for (int a = 0; a < 3; a++) {
                const double w = M[a * 4 + 0] * x + M[a * 4 + 1] * y + M[a * 4 + 2] * z + M[a * 4 + 3];
                mn[a] = std::min(mn[a], (float)w); mx[a] = std::max(mx[a], (float)w);
            }
        }
        for (int a = 0; a < 3; a++) {
            const float pad = 1e-5f * std::max(std::fabs(mn[a]), std::fabs(mx[a])) + 1e-6f;
            mn[a] -= pad; mx[a] += pad;
        }
    }

    // Children of inner node i: c = hit(i); while c != miss(i): take c, c = miss(c).
    // Emits nodes in depth-first order so that leaf numbers follow the reference's visit order.
    int32_t build(int32_t i, int32_t depth)
    {
        if (i < 0 || (uint32_t)i >= count) { out->error = "link out of range in node list"; return kRefNull; }
        if (depth > 96) { out->error = "node list deeper than 96 levels (cyclic links?)"; return kRefNull; }
        if (depth > depth_seen) depth_seen = depth;
        const AtnBvhNode& n = list[i];
        if (is_leaf(n)) return make_leaf(n);

        const int32_t escape = (int32_t)n.miss;
        int32_t kids[2]; int nk = 0;
        int32_t c = (int32_t)n.hit;
        while (c >= 0 && c != escape) {
            if (nk == 2) { out->error = "inner node with more than two children"; return kRefNull; }
            kids[nk++] = c;
            c = (int32_t)list[c].miss;   // inner child: its sibling; leaf child: hit == miss == next in preorder
        }
        if (nk == 0) return kRefNull;

        const int32_t self = (int32_t)out->nodes.size();
        out->nodes.emplace_back();
        int32_t refs[2] = { kRefNull, kRefNull };
        for (int k = 0; k < nk; k++) refs[k] = build(kids[k], depth + 1);
        PairNode& pn = out->nodes[self];
        // A missing second child keeps ref kRefNull; the kernels skip it without a box test.
        const float big = 3.0e38f;
        float lmin[3] = { big, big, big }, lmax[3] = { -big, -big, -big };
        float rmin[3] = { big, big, big }, rmax[3] = { -big, -big, -big };
        for (int a = 0; a < 3; a++) { lmin[a] = list[kids[0]].boxmin[a]; lmax[a] = list[kids[0]].boxmax[a]; }
        if (nk == 2) for (int a = 0; a < 3; a++) { rmin[a] = list[kids[1]].boxmin[a]; rmax[a] = list[kids[1]].boxmax[a]; }
        // The reference never reads a LEAF's box (threaded_bvh_traverser.h:142-220), and SBVH leaf boxes
        // are clipped pieces that need not contain the triangle.  The slot is re-used for a box that is
        // conservative by construction; the kernels use it only as a pre-filter (leaf_margin).
        if (refs[0] < 0 && refs[0] != kRefNull) leaf_bounds(refs[0], lmin, lmax);
        if (nk == 2 && refs[1] < 0 && refs[1] != kRefNull) leaf_bounds(refs[1], rmin, rmax);
        pn.n0[0] = lmin[0]; pn.n0[1] = lmax[0]; pn.n0[2] = lmin[1]; pn.n0[3] = lmax[1];
        pn.n1[0] = rmin[0]; pn.n1[1] = rmax[0]; pn.n1[2] = rmin[1]; pn.n1[3] = rmax[1];
        pn.n2[0] = lmin[2]; pn.n2[1] = lmax[2]; pn.n2[2] = rmin[2]; pn.n2[3] = rmax[2];
        pn.left = refs[0]; pn.right = refs[1]; pn.pad0 = 0; pn.pad1 = 0;
        return self;
    }
};

}  // namespace detail

// Root box of a list: the reference tests the root node's own box like any inner node
// (threaded_bvh_traverser.h:278-280).  A PairNode only stores child boxes, so every tree gets a
// one-child wrapper node whose left box is the root's box.
inline int32_t wrap_root(EncodedBvh* out, const AtnBvhNode& root, int32_t root_ref)
{
    if (root_ref < 0) return root_ref;   // a leaf root is never box-tested by the reference
    PairNode pn;
    const float big = 3.0e38f;
    pn.n0[0] = root.boxmin[0]; pn.n0[1] = root.boxmax[0]; pn.n0[2] = root.boxmin[1]; pn.n0[3] = root.boxmax[1];
    pn.n1[0] = big; pn.n1[1] = -big; pn.n1[2] = big; pn.n1[3] = -big;
    pn.n2[0] = root.boxmin[2]; pn.n2[1] = root.boxmax[2]; pn.n2[2] = big; pn.n2[3] = -big;
    pn.left = root_ref; pn.right = kRefNull; pn.pad0 = 0; pn.pad1 = 0;
    out->nodes.push_back(pn);
    return (int32_t)out->nodes.size() - 1;
}

inline EncodedBvh encode_bvh(const AtnSceneBlob* scene)
{
    EncodedBvh out;
    if (scene->node_list_count == 0) return out;
    std::vector<int32_t> roots(scene->node_list_count, kRefNull);
    for (uint32_t k = 1; k < scene->node_list_count; k++) {
        if (scene->node_list_sizes[k] == 0) continue;
        detail::ListBuilder lb{ scene->node_lists[k], scene->node_list_sizes[k], false, scene, &out, &roots };
        const int32_t r = lb.build(0, 0);
        roots[k] = wrap_root(&out, scene->node_lists[k][0], r);
        if (lb.depth_seen > out.max_depth) out.max_depth = lb.depth_seen;
        if (!out.error.empty()) return out;
    }
    if (scene->node_list_sizes[0] > 0) {
        detail::ListBuilder lb{ scene->node_lists[0], scene->node_list_sizes[0], true, scene, &out, &roots };
        const int32_t r = lb.build(0, 0);
        out.top_root = wrap_root(&out, scene->node_lists[0][0], r);
        out.max_depth += lb.depth_seen + 4;
    }
    return out;
}

}  // namespace atn
