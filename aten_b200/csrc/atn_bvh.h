// atn_bvh.h -- host-side re-encoding of the reference's threaded node lists into the traversal
// layout the sm_100a kernels walk (DESIGN.md "Data layout in HBM").
//
// Input  : std::vector<std::vector<GPUBvhNode>> exactly as the reference's builders emit them
//          (48-byte nodes, hit/miss links stored as floats; list 0 = top layer over instances,
//          list k = SBVH of object k, one triangle per leaf; reference
//          src/libaten/accelerator/threaded_bvh.h:13-56, sbvh.h:19-66, sbvh.cpp:827-950,
//          threaded_bvh.cpp:254-357).
// Output : * QuadNode[]  128 B, 128-byte aligned: up to four slots (box + reference) obtained by an
//            exact collapse of the binary tree (see "Wide (4-slot) encoding" below); one node fetch
//            (4 x LDG.256) decides four subtrees / leaves;
//          * TriLeaf[]   64 B: v0 and the two edges v1-v0, v2-v0 gathered next to each other with
//            primid / mtrlid / mesh_id in the .w lanes (no index -> vertex indirection);
//          * InstLeafW[] 64 B: objid, root reference of the bottom tree, W2L matrix index, meshid and a
//            conservative world-space box of the instance.
//          Leaves are numbered in the reference's depth-first (hit-link) order; the index is the
//          tie-break rank that reproduces "first encountered wins" (SURVEY Q2).
// Boxes stay fp32 and un-quantised: the reference's slab test (aabb.h:61-86, +1e-6 bias) is part of
// its result, and primitive ids must match bit for bit (see DESIGN.md for why quantisation was
// rejected for this round).
#pragma once

#include <stdint.h>
#include <string.h>
#include <algorithm>
#include <atomic>
#include <thread>
#include <cmath>
#include <limits>
#include <string>
#include <vector>

#include "../../include/aten_b200.h"

namespace atn {

struct alignas(32) TriLeaf {
    float v0[3]; int32_t primid;
    float e1[3]; int32_t mtrlid;      // v1 - v0 (the float subtraction intersectTriangle starts with)
    float e2[3]; int32_t mesh_id;     // v2 - v0
    int32_t pad[4];                   // 64 B: one 256-bit + one 128-bit load, 32-byte aligned
};
static_assert(sizeof(TriLeaf) == 64, "TriLeaf");

struct alignas(16) InstLeaf {
    int32_t objid;      // object index recorded in isect.objid
    int32_t root;       // reference of the bottom tree's root (see kRef*)
    int32_t w2l;        // index of the world-to-local matrix, -1: none (ray used as is)
    int32_t meshid;
};

// Child / root references: >= 0 inner QuadNode index; < 0 leaf number ~ref; two reserved values.
constexpr int32_t kRefNull = INT32_MIN + 1;   // leaf that holds nothing traversable (e.g. a sphere)
constexpr int32_t kRefExit = INT32_MIN;       // stack marker: leave the current instance
constexpr int32_t kRefDone = INT32_MIN + 2;   // stack sentinel below every ray's first entry: the traversal is complete

struct EncodedBvh {          // leaf records collected while a node list is scanned
    std::vector<TriLeaf> tris;
    std::vector<InstLeaf> insts;
    std::vector<int32_t> inst_list;   // per instance leaf: the number of its bottom node list
    std::string error;
};

// What the top layer needs to know about the encoded bottom trees.  A top-layer update (Renderer::updateBVH: list 0,
// objects and matrices change every frame, renderer.cpp:133-153) re-encodes list 0 against this and touches nothing else.
struct BottomInfo {
    std::vector<int32_t> roots;      // per node list: reference of its root, kRefNull for an empty list (entry 0 unused)
    std::vector<int32_t> need;       // per node list: stack entries the traversal can hold below that root
    std::vector<float> box;          // per node list: root box, min xyz then max xyz (node 0 of the reference's list)
    int32_t node_count{ 0 };         // QuadNodes of all bottom lists; the top layer's nodes follow them
};

inline int32_t leaf_ref(int32_t leaf_index) { return ~leaf_index; }

namespace detail {

struct ListBuilder {
    const AtnBvhNode* list;
    uint32_t count;
    bool top;
    const AtnSceneBlob* scene;
    EncodedBvh* out;
    const BottomInfo* bottom;                 // for the top layer: roots, stack needs and root boxes of the bottom lists

    static bool is_leaf(const AtnBvhNode& n) { return n.object_id >= 0 || n.primid >= 0; }

    int32_t make_leaf(const AtnBvhNode& n)
    {
        if (top) {
            if (!(n.exid >= 0)) return kRefNull;        // no external tree: sphere / cube, never hit (F5)
            uint32_t bits; memcpy(&bits, &n.exid, 4);
            const int32_t exid = (int32_t)(bits & 0x7fffu);   // mainExid:15 (threaded_bvh.h:29-37)
            if (exid <= 0 || exid >= (int32_t)bottom->roots.size()) return kRefNull;
            if ((uint32_t)n.object_id >= scene->object_count) { out->error = "object index out of range in the top layer"; return kRefNull; }
            InstLeaf il;
            il.objid = (int32_t)n.object_id;
            il.root = bottom->roots[exid];
            const AtnObjectParameter& obj = scene->objects[(uint32_t)n.object_id];
            il.w2l = obj.mtx_id >= 0 ? obj.mtx_id + 1 : -1;    // threaded_bvh_traverser.h:149-158
            il.meshid = (int32_t)n.meshid;
            out->insts.push_back(il);
            out->inst_list.push_back(exid);
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
        tl.pad[0] = tl.pad[1] = tl.pad[2] = tl.pad[3] = 0;
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
        // the instance's root box: node 0 of its bottom list
        const int32_t lk = out->inst_list[idx];
        const float* rb = (il.root != kRefNull && (size_t)lk * 6 + 5 < bottom->box.size()) ? &bottom->box[(size_t)lk * 6] : nullptr;
        const float big = 3.0e38f;
        if (!rb) { for (int a = 0; a < 3; a++) { mn[a] = -big; mx[a] = big; } return; }
        float M[16] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
        if (obj.mtx_id >= 0 && (uint32_t)obj.mtx_id < scene->matrix_count) memcpy(M, scene->matrices[obj.mtx_id].m, sizeof(M));
        for (int a = 0; a < 3; a++) { mn[a] = big; mx[a] = -big; }
        for (int c = 0; c < 8; c++) {
            const double x = (c & 1) ? rb[3] : rb[0];
            const double y = (c & 2) ? rb[4] : rb[1];
            const double z = (c & 4) ? rb[5] : rb[2];
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
};

}  // namespace detail

// ---------------------------------------------------------------------------------------------
// Wide (4-slot) encoding.
//
// The reference's trees are binary with one triangle per leaf.  A QuadNode stands for one inner node P
// of the reference's tree whose own box has already been tested, and holds up to four SLOTS obtained
// by replacing children of P by their own children (largest box first) while the result fits.  What makes
// this exact rather than approximate:
//   * an inner node g below a replaced child X keeps its own box.  The reference reaches g only if X's
//     box was hit as well; g's box lies inside X's box plane by plane (checked here, node by node --
//     X is only replaced if it holds), and the slab test aabb.h:61-86 is monotone in the plane values
//     (same multiplier, same addend, rounding is monotone), so hit(g) implies hit(X) bit for bit;
//   * a LEAF below a replaced child X is never box-tested by the reference; it is visited iff X's box
//     was hit.  Its slot therefore carries X's box;
//   * a leaf that is a child of P itself is visited whenever P is: its slot carries an always-hit box.
// Slots keep the reference's depth-first order and leaves keep their depth-first numbers, so the
// tie-break rank is unchanged.
// ---------------------------------------------------------------------------------------------
struct alignas(128) QuadNode {
    // slots (0,1): q[0] = min.x0, min.y0, min.x1, min.y1 ; q[1] = max.x0, max.y0, max.x1, max.y1 ;
    //              q[2] = min.z0, min.z1, max.z0, max.z1
    // slots (2,3): q[3], q[4], q[5] likewise.  Pairs (x, y) share the (invdir.x, invdir.y) multiplier pair.
    float q[6][4];
    int32_t ref[4];        // >= 0 QuadNode index, < 0 leaf ~ref, kRefNull: empty slot
    int32_t pad[4];
};
static_assert(sizeof(QuadNode) == 128, "QuadNode");

struct alignas(32) InstLeafW {
    int32_t objid, root, w2l, meshid;
    float wmin[3]; float pad0;      // conservative world-space bounds of the instance (entry pre-filter)
    float wmax[3]; float pad1;
    int32_t pad2[4];
};
static_assert(sizeof(InstLeafW) == 64, "InstLeafW");

struct WideBvh {
    std::vector<QuadNode> nodes;
    std::vector<TriLeaf> tris;
    std::vector<InstLeafW> insts;
    int32_t top_root{ kRefNull };
    int32_t max_stack{ 0 };          // exact worst case of the traversal stack depth (entries), see stack_need
    std::vector<int32_t> stack_need; // per QuadNode: entries the traversal can have on its stack below this node
    uint64_t slots_used{ 0 }, replaced{ 0 }, not_nested{ 0 };
    std::string error;
};

namespace detail {

struct WideBuilder {
    const AtnBvhNode* list;
    uint32_t count;
    bool top;
    EncodedBvh* leaves;                       // receives TriLeaf / InstLeaf records in depth-first order
    ListBuilder* lb;                          // leaf construction and bounds
    WideBvh* out;
    std::vector<int32_t> kid0, kid1, leafref; // per original node
    std::vector<uint8_t> nested;
    int32_t depth_seen{ 0 };

    bool inner(int32_t i) const { return !ListBuilder::is_leaf(list[i]); }

    // pass 1: children and leaves in the reference's depth-first order
    void scan(int32_t i, int32_t depth)
    {
        if (!out->error.empty()) return;
        if (i < 0 || (uint32_t)i >= count) { out->error = "link out of range in node list"; return; }
        if (depth > 96) { out->error = "node list deeper than 96 levels (cyclic links?)"; return; }
        const AtnBvhNode& n = list[i];
        if (ListBuilder::is_leaf(n)) {
            leafref[i] = lb->make_leaf(n);
            if (!leaves->error.empty()) out->error = leaves->error;
            return;
        }
        const int32_t escape = (int32_t)n.miss;
        int nk = 0;
        int32_t c = (int32_t)n.hit;
        while (c >= 0 && c != escape) {
            if ((uint32_t)c >= count) { out->error = "link out of range in node list"; return; }
            if (nk == 2) { out->error = "inner node with more than two children"; return; }
            (nk == 0 ? kid0 : kid1)[i] = c;
            nk++;
            c = (int32_t)list[c].miss;
        }
        bool ok = true;
        const int32_t ks[2] = { kid0[i], kid1[i] };
        for (int k = 0; k < nk; k++) {
            scan(ks[k], depth + 1);
            if (!out->error.empty()) return;
            if (inner(ks[k])) {
                const AtnBvhNode& g = list[ks[k]];
                for (int a = 0; a < 3; a++)
                    if (!(g.boxmin[a] >= n.boxmin[a] && g.boxmax[a] <= n.boxmax[a])) ok = false;
            }
        }
        nested[i] = ok ? 1 : 0;
        if (!ok) out->not_nested++;
    }

    struct Slot { int32_t node, boxnode; };   // boxnode: original node whose box the slot tests, -1 = always hit
    struct SlotList {                         // at most four slots, plus the two a replacement inserts before one is dropped
        Slot s[6];
        int n{ 0 };
    };

    static float area(const AtnBvhNode& n)
    {
        const float dx = n.boxmax[0] - n.boxmin[0], dy = n.boxmax[1] - n.boxmin[1], dz = n.boxmax[2] - n.boxmin[2];
        return dx * dy + dy * dz + dz * dx;
    }

    // children of X in order, inserted at `at`; a leaf child inherits X's box unless X's box is known to be hit
    void push_kids(SlotList& slots, int at, int32_t X, bool x_tested)
    {
        Slot ins[2];
        int ni = 0;
        const int32_t ks[2] = { kid0[X], kid1[X] };
        for (int k = 0; k < 2; k++) {
            if (ks[k] < 0) continue;
            if (inner(ks[k])) ins[ni++] = Slot{ ks[k], ks[k] };
            else if (leafref[ks[k]] != kRefNull) ins[ni++] = Slot{ ks[k], x_tested ? -1 : X };
        }
        for (int k = slots.n - 1; k >= at; k--) slots.s[k + ni] = slots.s[k];
        for (int k = 0; k < ni; k++) slots.s[at + k] = ins[k];
        slots.n += ni;
    }

    int nkids_live(int32_t X) const
    {
        int n = 0;
        const int32_t ks[2] = { kid0[X], kid1[X] };
        for (int k = 0; k < 2; k++)
            if (ks[k] >= 0 && (inner(ks[k]) || leafref[ks[k]] != kRefNull)) n++;
        return n;
    }

    // stack entries the traversal can hold while it is below reference `ref`: an instance leaves an exit marker
    // behind and then walks the bottom tree (emitted before the top layer)
    int32_t need_of(int32_t ref) const
    {
        if (ref >= 0) return (size_t)ref < out->stack_need.size() ? out->stack_need[ref] : 0;
        if (ref == kRefNull || !top) return 0;
        const int32_t lk = leaves->inst_list[~ref];
        return 1 + lb->bottom->need[lk];
    }

    // QuadNode for the slots `slots` (already the children of a tested node); returns its index
    int32_t emit(SlotList slots, int32_t depth)
    {
        if (depth > depth_seen) depth_seen = depth;
        for (;;) {
            int best = -1; float best_area = -1.0f;
            for (int s = 0; s < slots.n; s++) {
                const int32_t X = slots.s[s].node;
                if (!inner(X) || !nested[X]) continue;
                if (slots.n - 1 + nkids_live(X) > 4) continue;
                const float a = area(list[X]);
                if (a > best_area) { best_area = a; best = (int)s; }
            }
            if (best < 0) break;
            const int32_t X = slots.s[best].node;
            for (int k = best; k + 1 < slots.n; k++) slots.s[k] = slots.s[k + 1];
            slots.n--;
            push_kids(slots, best, X, false);
            out->replaced++;
        }
        if (slots.n == 0) return kRefNull;
        const int32_t self = (int32_t)out->nodes.size();
        out->nodes.emplace_back();
        int32_t refs[4] = { kRefNull, kRefNull, kRefNull, kRefNull };
        for (int s = 0; s < slots.n; s++) {
            const int32_t X = slots.s[s].node;
            if (inner(X)) {
                SlotList sub;
                push_kids(sub, 0, X, true);
                refs[s] = emit(sub, depth + 1);
            }
            else refs[s] = leafref[X];
        }
        QuadNode& qn = out->nodes[self];
        const float big = 3.0e38f;
        for (int s = 0; s < 4; s++) {
            // empty slot: min = max = +inf.  Its slab distances are +inf or -inf on every axis, so entry <= min(exit, t_max)
            // fails for any finite t_max and the kernel needs no separate test of the reference
            const float inf = std::numeric_limits<float>::infinity();
            float mn[3] = { inf, inf, inf }, mx[3] = { inf, inf, inf };
            if (s < slots.n && refs[s] != kRefNull) {
                if (slots.s[s].boxnode < 0) { for (int a = 0; a < 3; a++) { mn[a] = -big; mx[a] = big; } }
                else for (int a = 0; a < 3; a++) { mn[a] = list[slots.s[s].boxnode].boxmin[a]; mx[a] = list[slots.s[s].boxnode].boxmax[a]; }
                out->slots_used++;
            }
            const int g = (s >> 1) * 3, h = (s & 1);
            qn.q[g + 0][h * 2 + 0] = mn[0]; qn.q[g + 0][h * 2 + 1] = mn[1];
            qn.q[g + 1][h * 2 + 0] = mx[0]; qn.q[g + 1][h * 2 + 1] = mx[1];
            qn.q[g + 2][h] = mn[2]; qn.q[g + 2][2 + h] = mx[2];
            qn.ref[s] = refs[s];
            qn.pad[s] = 0;
        }
        // Worst case below this node: the deepest child is entered first, with all other hit slots left on the stack.
        int32_t live = 0, deepest = 0;
        for (int s = 0; s < 4; s++) {
            if (refs[s] == kRefNull) continue;
            live++;
            deepest = std::max(deepest, need_of(refs[s]));
        }
        if (out->stack_need.size() < out->nodes.size()) out->stack_need.resize(out->nodes.size(), 0);
        out->stack_need[self] = (live > 0 ? live - 1 : 0) + deepest;
        return self;
    }

    // the whole list: the root's own box is tested like any inner node's (threaded_bvh_traverser.h:278-280)
    int32_t run()
    {
        kid0.assign(count, -1); kid1.assign(count, -1); leafref.assign(count, kRefNull); nested.assign(count, 0);
        scan(0, 0);
        if (!out->error.empty()) return kRefNull;
        if (!inner(0)) return leafref[0];        // a leaf root is never box-tested by the reference
        SlotList slots;
        slots.s[0] = Slot{ 0, 0 }; slots.n = 1;
        return emit(slots, 1);
    }
};

}  // namespace detail

// The top layer alone (node list 0) against already encoded bottom trees: its QuadNodes (references numbered from
// bottom.node_count on), the instance leaves with their world boxes, the root and the exact worst-case stack depth of
// the whole two-level walk.  `scene` supplies objects and matrices only.
inline WideBvh encode_top(const AtnBvhNode* top_nodes, uint32_t top_count, const AtnSceneBlob* scene, const BottomInfo& bottom)
{
    WideBvh out;
    if (top_count == 0) return out;
    EncodedBvh leaves;
    detail::ListBuilder lb{ top_nodes, top_count, true, scene, &leaves, &bottom };
    detail::WideBuilder wb{ top_nodes, top_count, true, &leaves, &lb, &out };
    out.top_root = wb.run();
    if (out.error.empty() && !leaves.error.empty()) out.error = leaves.error;
    if (!out.error.empty()) return out;
    const int32_t node_off = bottom.node_count;
    for (QuadNode& qn : out.nodes)
        for (int sl = 0; sl < 4; sl++) if (qn.ref[sl] >= 0) qn.ref[sl] += node_off;     // leaf references are instance numbers
    if (out.top_root >= 0) out.top_root += node_off;
    out.insts.resize(leaves.insts.size());
    for (size_t i = 0; i < leaves.insts.size(); i++) {
        const InstLeaf& il = leaves.insts[i];
        InstLeafW& w = out.insts[i];
        w.objid = il.objid; w.root = il.root; w.w2l = il.w2l; w.meshid = il.meshid;
        float mn[3], mx[3];
        lb.leaf_bounds(leaf_ref((int32_t)i), mn, mx);
        for (int a = 0; a < 3; a++) { w.wmin[a] = mn[a]; w.wmax[a] = mx[a]; }
        w.pad0 = 0.0f; w.pad1 = 0.0f; w.pad2[0] = w.pad2[1] = w.pad2[2] = w.pad2[3] = 0;
    }
    out.max_stack = 0;
    if (out.top_root >= 0) out.max_stack = out.stack_need[out.top_root - node_off];
    else if (out.top_root != kRefNull && !out.insts.empty())       // the top layer is a single instance leaf
        out.max_stack = 1 + bottom.need[leaves.inst_list[~out.top_root]];
    out.max_stack += 2;    // the sentinel at the bottom of every ray's stack, and one entry of slack
    return out;
}

// The encoded bottom trees of a scene as a unit: what an encoded-tree file holds (atn_cache.h).
struct CachedBottom {
    BottomInfo info;
    std::vector<QuadNode> nodes;
    std::vector<int32_t> stack_need;
    std::vector<TriLeaf> tris;
    uint64_t slots_used{ 0 }, replaced{ 0 }, not_nested{ 0 };
};

// The whole scene.  `info` (optional) receives what later top-layer updates need (encode_top).  `cached` (optional): the
// bottom trees as an earlier encode_wide of the same node lists produced them -- only the top layer is encoded then.
inline WideBvh encode_wide(const AtnSceneBlob* scene, BottomInfo* info = nullptr, const CachedBottom* cached = nullptr)
{
    WideBvh out;
    if (scene->node_list_count == 0) return out;
    const uint32_t nlists = scene->node_list_count;
    BottomInfo bottom;
    if (cached) {
        out.nodes = cached->nodes; out.stack_need = cached->stack_need; out.tris = cached->tris;
        out.slots_used = cached->slots_used; out.replaced = cached->replaced; out.not_nested = cached->not_nested;
        bottom = cached->info;
        if (bottom.roots.size() != nlists) { out.error = "cached bottom trees belong to another scene"; return out; }
    }
    else {
    bottom.roots.assign(nlists, kRefNull); bottom.need.assign(nlists, 0); bottom.box.assign((size_t)nlists * 6, 0.0f);
    // Bottom lists are independent of each other: each is encoded on its own (in parallel when the scene is large) with
    // list-local node and leaf numbers, then appended in list order with the numbers moved to their global positions --
    // the result is what the lists encoded one after the other into shared arrays would give.
    struct Part { WideBvh w; EncodedBvh l; int32_t root{ kRefNull }; };
    std::vector<Part> parts(nlists);
    auto encode_list = [&](uint32_t k) {
        Part& pt = parts[k];
        const uint32_t cnt = scene->node_list_sizes[k];
        pt.l.tris.reserve(cnt / 2 + 1); pt.w.nodes.reserve(cnt / 4 + 1); pt.w.stack_need.reserve(cnt / 4 + 1);
        detail::ListBuilder lb{ scene->node_lists[k], cnt, false, scene, &pt.l, nullptr };
        detail::WideBuilder wb{ scene->node_lists[k], cnt, false, &pt.l, &lb, &pt.w };
        pt.root = wb.run();
    };
    size_t total = 0;
    for (uint32_t k = 1; k < nlists; k++) total += scene->node_list_sizes[k];
    {
        unsigned nthreads = std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
        if (total < 400000 || nlists < 3) nthreads = 1;
        std::atomic<uint32_t> next{ 1 };
        auto worker = [&]() {
            for (uint32_t k = next.fetch_add(1); k < nlists; k = next.fetch_add(1))
                if (scene->node_list_sizes[k] > 0) encode_list(k);
        };
        std::vector<std::thread> pool;
        for (unsigned t = 1; t < nthreads; t++) pool.emplace_back(worker);
        worker();
        for (auto& th : pool) th.join();
    }
    out.nodes.reserve(total / 4 + scene->node_list_sizes[0] + 1);
    out.tris.reserve(total / 2 + 1);
    for (uint32_t k = 1; k < nlists; k++) {
        if (scene->node_list_sizes[k] == 0) continue;
        Part& pt = parts[k];
        if (!pt.w.error.empty()) { out.error = pt.w.error; return out; }
        const int32_t node_off = (int32_t)out.nodes.size(), leaf_off = (int32_t)out.tris.size();
        auto move_ref = [&](int32_t ref) -> int32_t {
            if (ref >= 0) return ref + node_off;
            if (ref == kRefNull || ref == kRefExit || ref == kRefDone) return ref;
            return leaf_ref(~ref + leaf_off);
        };
        for (QuadNode& qn : pt.w.nodes)
            for (int sl = 0; sl < 4; sl++) qn.ref[sl] = move_ref(qn.ref[sl]);
        out.nodes.insert(out.nodes.end(), pt.w.nodes.begin(), pt.w.nodes.end());
        out.stack_need.insert(out.stack_need.end(), pt.w.stack_need.begin(), pt.w.stack_need.end());
        out.tris.insert(out.tris.end(), pt.l.tris.begin(), pt.l.tris.end());
        out.slots_used += pt.w.slots_used; out.replaced += pt.w.replaced; out.not_nested += pt.w.not_nested;
        bottom.roots[k] = move_ref(pt.root);
        bottom.need[k] = pt.root >= 0 ? pt.w.stack_need[pt.root] : 0;
        const AtnBvhNode& n0 = scene->node_lists[k][0];
        for (int a = 0; a < 3; a++) { bottom.box[(size_t)k * 6 + a] = n0.boxmin[a]; bottom.box[(size_t)k * 6 + 3 + a] = n0.boxmax[a]; }
        pt = Part();                                  // release the list's partial arrays
    }
    bottom.node_count = (int32_t)out.nodes.size();
    }       // !cached
    if (scene->node_list_sizes[0] > 0) {
        WideBvh top = encode_top(scene->node_lists[0], scene->node_list_sizes[0], scene, bottom);
        if (!top.error.empty()) { out.error = top.error; return out; }
        out.nodes.insert(out.nodes.end(), top.nodes.begin(), top.nodes.end());
        out.stack_need.insert(out.stack_need.end(), top.stack_need.begin(), top.stack_need.end());
        out.insts.swap(top.insts);
        out.top_root = top.top_root;
        out.max_stack = top.max_stack;
        out.slots_used += top.slots_used; out.replaced += top.replaced; out.not_nested += top.not_nested;
    }
    else out.max_stack = 2;
    if (info) *info = std::move(bottom);
    return out;
}

}  // namespace atn
