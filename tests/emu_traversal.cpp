// TEST INFRASTRUCTURE (not part of the product, not loaded by it): a sequential CPU emulation of the
// traversal kernel's per-ray logic (k_trace, atn_kernels.cu) over the QuadNode encoding that the
// product's host code builds (atn_bvh.h, encode_wide).  It lets the CPU suite check the encoder -- the
// exact collapse of the reference's binary trees -- against the oracle without a GPU, and counts node
// steps / triangle tests / stack depth (scripts/emu_wide.py).
//   g++ -std=c++17 -O2 -fopenmp -fPIC -ffp-contract=off -shared -o tests/_build/libemu_traversal.so tests/emu_traversal.cpp
#include <stdint.h>
#include <string.h>
#include <cmath>
#include <vector>

#include "../aten_b200/csrc/atn_math.cuh"
#include "../aten_b200/csrc/atn_bvh.h"
#include "../aten_b200/csrc/atn_cache.h"

using namespace atn;

namespace {

constexpr float kBig = 3.402823466e+38f;

struct Slab { float t0, t1; bool hit; };
inline Slab slab(const RayInv& ri, const float mn[3], const float mx[3], float t_min, float t_max)
{
    const float fx = mx[0] * ri.invdir.x + ri.oxinvdir.x;
    const float fy = mx[1] * ri.invdir.y + ri.oxinvdir.y;
    const float fz = mx[2] * ri.invdir.z + ri.oxinvdir.z;
    const float nx = mn[0] * ri.invdir.x + ri.oxinvdir.x;
    const float ny = mn[1] * ri.invdir.y + ri.oxinvdir.y;
    const float nz = mn[2] * ri.invdir.z + ri.oxinvdir.z;
    Slab s;
    s.t1 = fminf(fminf(fmaxf(fx, nx), fmaxf(fy, ny)), fmaxf(fz, nz));
    s.t0 = fmaxf(fmaxf(fmaxf(fminf(fx, nx), fminf(fy, ny)), fminf(fz, nz)), t_min);
    s.hit = s.t0 <= fminf(s.t1, t_max);
    return s;
}

inline float leaf_margin(const V3& org, const V3& dir, const RayInv& ri)
{
    const float dmin = fminf(fminf(fabsf(dir.x), fabsf(dir.y)), fabsf(dir.z));
    if (!(dmin >= 0.02f)) return -1.0f;
    const float S = fabsf(ri.oxinvdir.x) + fabsf(ri.oxinvdir.y) + fabsf(ri.oxinvdir.z);
    const float M = fmaxf(fmaxf(fabsf(org.x), fabsf(org.y)), fabsf(org.z));
    return 1e-6f * S + 5e-4f * (1.0f + M);
}
inline bool leaf_may_hit(float t0, float t1, float margin)
{
    return (t0 <= t1 + 2e-4f * (fabsf(t0) + fabsf(t1)) + margin) | (margin < 0.0f);
}

inline TriHit tri_edges(const V3& org, const V3& dir, const V3& v0, const V3& e1, const V3& e2)
{
    const V3 r = org - v0;
    const V3 u = cross(dir, e2);
    const V3 v = cross(r, e1);
    const float inv = 1.0f / dot(u, e1);
    TriHit h;
    h.t = dot(v, e2) * inv;
    h.a = dot(u, r) * inv;
    h.b = dot(v, dir) * inv;
    h.hit = ((h.a >= 0.0f && h.a <= 1.0f) && (h.b >= 0.0f && h.b <= 1.0f) && (h.a + h.b <= 1.0f) && h.t >= 0.0f);
    return h;
}

struct Best { float t, a, b; int32_t leaf, inst; };
struct Ent { int32_t ref; float t0; };

struct Stats { uint64_t nodes, tris, entries, entries_filtered, pops_culled, rays; uint64_t max_sp; uint64_t sum_sp; uint64_t slot_hist[5]; };

inline void finish(const std::vector<TriLeaf>& tris, const int32_t* inst_objid, const int32_t* inst_meshid, const Best& b, AtnHit* h)
{
    h->t = b.t; h->a = b.a; h->b = b.b; h->isVoxel = 0;
    if (b.leaf < 0) { h->objid = -1; h->primid = -1; h->mtrlid = -1; h->meshid = -1; return; }
    const TriLeaf& t = tris[b.leaf];
    h->primid = t.primid; h->mtrlid = t.mtrlid;
    h->objid = inst_objid[b.inst];
    h->meshid = t.mesh_id < 0 ? inst_meshid[b.inst] : t.mesh_id;
}

inline float cull_bound(float t) { return t + fabsf(t) * 2e-5f; }

}  // namespace

// `band` (optional, 4 floats: rel, inv, S, M) switches on the near-tie bookkeeping of the kernel: a per-ray band
//   B(t) = |t| * rho + alpha,  rho = max(rel, inv * 1e-6 * max_k |invdir_k|),  alpha = S * sum_k |oxinvdir_k| + M * (1 + max_k |org_k|)
// (maxima over the world ray and every instance-space ray entered so far) bounds how far the biased slab entry distance of a
// box can lie beyond the hit distance of a triangle inside it.  Boxes are culled with t + 2 B(t); the smallest hit distance
// of a primitive other than the best one is tracked, and a ray whose runner-up lies within B of its best hit is FLAGGED:
// the reference's answer for it depends on the reference's visit order, and the exact (reference-order) walk decides it --
// unless the best hit is strictly the nearest and every box on the path to its leaf starts no later than the runner-up
// (stack keys carry the largest entry distance on the path, "pathmax"): the reference then reaches and keeps it whatever
// its order.
static int trace_impl(const AtnSceneBlob* scene, const AtnRay* rays, const float* t_max_in, size_t n, AtnHit* hits,
                      uint64_t* stats_out, int sort_leaves_first, int entry_filter, const float* band, uint8_t* flags)
{
    WideBvh w = encode_wide(scene);
    if (!w.error.empty()) return -1;
    std::vector<int32_t> iobj(w.insts.size()), imesh(w.insts.size());
    for (size_t i = 0; i < w.insts.size(); i++) { iobj[i] = w.insts[i].objid; imesh[i] = w.insts[i].meshid; }
    const float t_min = scene->config.bvh_hit_min > 0 ? scene->config.bvh_hit_min : kEps;
    Stats tot{};
#pragma omp parallel
    {
        Stats st{};
        std::vector<Ent> stack;
#pragma omp for schedule(dynamic, 1024)
        for (long long i = 0; i < (long long)n; i++) {
            const V3 worg(rays[i].org), wdir(rays[i].dir);
            V3 org = worg, dir = wdir;
            RayInv ri = ray_inv(org, dir);
            const RayInv wri = ri;
            const float wmargin = leaf_margin(org, dir, ri);
            float t_cull = t_max_in ? t_max_in[i] : kBig;
            Best best{ kBig, 0, 0, -1, -1 };
            float rho = 0.0f, alpha = 0.0f, t2 = kBig;
            auto widen = [&](const V3& o, const RayInv& r_) {
                if (!band) return;
                const float im = fmaxf(fmaxf(fabsf(r_.invdir.x), fabsf(r_.invdir.y)), fabsf(r_.invdir.z));
                const float S = fabsf(r_.oxinvdir.x) + fabsf(r_.oxinvdir.y) + fabsf(r_.oxinvdir.z);
                const float M = fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z));
                rho = fmaxf(rho, fmaxf(band[0], band[1] * 1e-6f * im));
                alpha = fmaxf(alpha, band[2] * S + band[3] * (1.0f + M));
            };
            auto cull_of = [&](float t) { return band ? t + 2.0f * (fabsf(t) * rho + alpha) : cull_bound(t); };
            widen(org, ri);
            stack.clear();
            int32_t cur = w.top_root, cur_inst = -1;
            float cur_key = -kBig, pm_best = kBig;
            for (;;) {
                if (cur >= 0) {
                    const QuadNode& q = w.nodes[cur];
                    st.nodes++;
                    { int used = 0; for (int k = 0; k < 4; k++) used += q.ref[k] != kRefNull; st.slot_hist[used]++; }
                    Ent e[4]; int ne = 0;
                    for (int s = 0; s < 4; s++) {
                        if (q.ref[s] == kRefNull) continue;
                        const int g = (s >> 1) * 3, h = s & 1;
                        const float mn[3] = { q.q[g][h * 2], q.q[g][h * 2 + 1], q.q[g + 2][h] };
                        const float mx[3] = { q.q[g + 1][h * 2], q.q[g + 1][h * 2 + 1], q.q[g + 2][2 + h] };
                        const Slab sl = slab(ri, mn, mx, t_min, t_cull);
                        if (!sl.hit) continue;
                        e[ne].ref = q.ref[s];
                        e[ne].t0 = (sort_leaves_first && q.ref[s] < 0) ? -kBig : (band ? fmaxf(sl.t0, cur_key) : sl.t0);
                        ne++;
                    }
                    // stable insertion sort by entry distance
                    for (int a = 1; a < ne; a++) { Ent k = e[a]; int b = a - 1; while (b >= 0 && e[b].t0 > k.t0) { e[b + 1] = e[b]; b--; } e[b + 1] = k; }
                    for (int a = ne - 1; a >= 1; a--) stack.push_back(e[a]);
                    cur = ne ? e[0].ref : kRefNull;
                    if (ne) cur_key = e[0].t0;
                }
                else if (cur != kRefNull && cur != kRefExit) {
                    const int32_t leaf = ~cur;
                    if (cur_inst >= 0) {
                        const TriLeaf& t = w.tris[leaf];
                        st.tris++;
                        const TriHit h = tri_edges(org, dir, V3(t.v0[0], t.v0[1], t.v0[2]), V3(t.e1[0], t.e1[1], t.e1[2]), V3(t.e2[0], t.e2[1], t.e2[2]));
                        if (h.hit && h.t < kBig && t_min < h.t) {
                            const bool closer = h.t < best.t || (h.t == best.t && (cur_inst < best.inst || (cur_inst == best.inst && leaf < best.leaf)));
                            const bool same_prim = best.leaf >= 0 && h.t == best.t && h.a == best.a && h.b == best.b;   // a duplicate SBVH leaf of the best primitive
                            if (closer) {
                                if (!same_prim && best.leaf >= 0) t2 = fminf(t2, best.t);
                                best.t = h.t; best.a = h.a; best.b = h.b; best.leaf = leaf; best.inst = cur_inst; t_cull = cull_of(h.t);
                                pm_best = same_prim ? fminf(pm_best, cur_key) : cur_key;
                            }
                            else if (!same_prim) t2 = fminf(t2, h.t);
                            else pm_best = fminf(pm_best, cur_key);
                        }
                        cur = kRefNull;
                    }
                    else {
                        const InstLeafW& il = w.insts[leaf];
                        cur = kRefNull;
                        bool enter = il.root != kRefNull;
                        if (enter && entry_filter) {
                            const Slab sl = slab(ri, il.wmin, il.wmax, t_min, kBig);
                            const float far = best.leaf >= 0 ? t_cull : kBig;
                            if (!leaf_may_hit(sl.t0, fminf(sl.t1, far), wmargin)) { enter = false; st.entries_filtered++; }
                        }
                        if (enter) {
                            st.entries++;
                            stack.push_back(Ent{ kRefExit, -kBig });
                            cur_inst = leaf;
                            if (il.w2l >= 0) {
                                const AtnMat4& m = scene->matrices[il.w2l];
                                const Ray tr = make_ray(mat_apply(m, org), mat_apply_xyz(m, dir));
                                org = tr.org; dir = tr.dir; ri = ray_inv(org, dir);
                                widen(org, ri);
                                if (band && best.leaf >= 0) t_cull = cull_of(best.t);
                            }
                            cur = il.root;
                        }
                    }
                }
                if (stack.size() > st.max_sp) st.max_sp = stack.size();
                bool finished = false;
                while (cur == kRefNull) {
                    if (stack.empty()) { finished = true; break; }
                    const Ent e = stack.back(); stack.pop_back();
                    cur = e.ref; cur_key = e.t0;
                    if (cur == kRefExit) {
                        if (stack.empty()) { finished = true; break; }
                        cur_inst = -1; org = worg; dir = wdir; ri = wri; cur = kRefNull; continue;
                    }
                    if (e.t0 > t_cull) { cur = kRefNull; st.pops_culled++; continue; }
                }
                if (finished) break;
            }
            finish(w.tris, iobj.data(), imesh.data(), best, &hits[i]);
            const bool certain = best.t < t2 && pm_best <= t2;
            if (flags) flags[i] = (band && best.leaf >= 0 && !certain && t2 <= best.t + (fabsf(best.t) * rho + alpha)) ? 1 : 0;
            st.rays++;
        }
#pragma omp critical
        {
            tot.nodes += st.nodes; tot.tris += st.tris; tot.entries += st.entries; tot.entries_filtered += st.entries_filtered;
            tot.pops_culled += st.pops_culled; tot.rays += st.rays; if (st.max_sp > tot.max_sp) tot.max_sp = st.max_sp;
            for (int k = 0; k < 5; k++) tot.slot_hist[k] += st.slot_hist[k];
        }
    }
    if (stats_out) {
        for (int k = 0; k < 5; k++) stats_out[11 + k] = tot.slot_hist[k];
        stats_out[0] = tot.nodes; stats_out[1] = tot.tris; stats_out[2] = tot.entries; stats_out[3] = tot.entries_filtered;
        stats_out[4] = tot.pops_culled; stats_out[5] = tot.max_sp; stats_out[6] = w.nodes.size(); stats_out[7] = w.slots_used;
        stats_out[8] = w.replaced; stats_out[9] = w.not_nested; stats_out[10] = (uint64_t)w.max_stack;
    }
    return 0;
}

extern "C" int emu_trace_wide(const AtnSceneBlob* scene, const AtnRay* rays, const float* t_max_in, size_t n, AtnHit* hits,
                              uint64_t* stats_out, int sort_leaves_first, int entry_filter)
{
    return trace_impl(scene, rays, t_max_in, n, hits, stats_out, sort_leaves_first, entry_filter, nullptr, nullptr);
}
extern "C" int emu_trace_wide_flag(const AtnSceneBlob* scene, const AtnRay* rays, const float* t_max_in, size_t n, AtnHit* hits,
                                   uint64_t* stats_out, const float* band, uint8_t* flags)
{
    return trace_impl(scene, rays, t_max_in, n, hits, stats_out, 0, 1, band, flags);
}

// Top-layer update (Renderer::updateBVH): encode_top(list 0 of `moved`) against the bottom trees kept from `base` must
// give exactly what a full encode of `moved` gives behind the bottom trees.  Returns 0 when identical, else a code.
extern "C" int emu_top_update_equals_fresh(const AtnSceneBlob* base, const AtnSceneBlob* moved)
{
    BottomInfo bi;
    const WideBvh a = encode_wide(base, &bi);
    if (!a.error.empty()) return -1;
    const WideBvh t = encode_top(moved->node_lists[0], moved->node_list_sizes[0], moved, bi);
    const WideBvh f = encode_wide(moved);
    if (!t.error.empty() || !f.error.empty()) return -2;
    const size_t nb = (size_t)bi.node_count;
    if (f.nodes.size() != nb + t.nodes.size()) return 1;
    if (nb && memcmp(a.nodes.data(), f.nodes.data(), nb * sizeof(QuadNode)) != 0) return 2;          // bottom trees untouched
    if (!t.nodes.empty() && memcmp(t.nodes.data(), f.nodes.data() + nb, t.nodes.size() * sizeof(QuadNode)) != 0) return 3;
    if (t.insts.size() != f.insts.size() || (!t.insts.empty() && memcmp(t.insts.data(), f.insts.data(), t.insts.size() * sizeof(InstLeafW)) != 0)) return 4;
    if (t.top_root != f.top_root) return 5;
    if (t.max_stack != f.max_stack) return 6;
    return 0;
}

// Encoded-tree cache (atn_cache.h): encode, write, read back, encode the top layer over the cached bottom trees -- the
// result must be byte-identical to the fresh encoding.  Returns 0 when it is; `other` (optional): a scene whose hash
// must NOT match the file.
extern "C" int emu_encoded_cache_round_trip(const AtnSceneBlob* scene, const char* path, const AtnSceneBlob* other)
{
    BottomInfo bi;
    const WideBvh fresh = encode_wide(scene, &bi);
    if (!fresh.error.empty()) return -1;
    const uint64_t hash = source_hash(scene);
    if (source_hash(scene) != hash) return -2;
    const size_t nb = (size_t)bi.node_count;
    if (!write_encoded(path, hash, scene->node_list_count, bi, fresh.nodes.data(), fresh.stack_need.data(), nb, fresh.tris.data(), fresh.tris.size(),
                       fresh.slots_used, fresh.replaced, fresh.not_nested).empty()) return -3;
    CachedBottom cb;
    if (!read_encoded(path, hash, scene->node_list_count, &cb).empty()) return -4;
    BottomInfo bi2;
    const WideBvh again = encode_wide(scene, &bi2, &cb);
    if (!again.error.empty()) return -5;
    if (again.nodes.size() != fresh.nodes.size() || memcmp(again.nodes.data(), fresh.nodes.data(), fresh.nodes.size() * sizeof(QuadNode)) != 0) return 1;
    if (again.tris.size() != fresh.tris.size() || memcmp(again.tris.data(), fresh.tris.data(), fresh.tris.size() * sizeof(TriLeaf)) != 0) return 2;
    if (again.insts.size() != fresh.insts.size() || (!fresh.insts.empty() && memcmp(again.insts.data(), fresh.insts.data(), fresh.insts.size() * sizeof(InstLeafW)) != 0)) return 3;
    if (again.top_root != fresh.top_root || again.max_stack != fresh.max_stack || again.stack_need != fresh.stack_need) return 4;
    if (bi2.roots != bi.roots || bi2.need != bi.need || bi2.box != bi.box || bi2.node_count != bi.node_count) return 5;
    if (other) {
        CachedBottom cb2;
        if (read_encoded(path, source_hash(other), other->node_list_count, &cb2).empty()) return 6;     // must be refused
    }
    if (read_encoded(path, hash ^ 1, scene->node_list_count, &cb).empty()) return 7;
    return 0;
}
