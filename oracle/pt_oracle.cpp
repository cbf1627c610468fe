// oracle/pt_oracle.cpp -- TEST INFRASTRUCTURE.  CPU restatement of the reference's path-tracing
// hot path over the flat scene blob (include/aten_b200.h).  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline leg may load this library; the product (libaten_b200.so) never
// links or calls it and has no CPU fallback.
//
// What is restated here, and from where:
//   * closest-hit traversal on the REFERENCE node layout and in the reference's visit order:
//     ThreadedBvhTraverser<true>::Traverse<Closest>
//     (src/libaten/accelerator/threaded_bvh_traverser.h:98-304), triangle::hit
//     (src/libaten/geometry/triangle.h:40-67);
//   * the per-pixel loop aten::PathTracing::radiance / OnRender
//     (src/libaten/renderer/pathtracing/pathtracing.cpp:23-89, 269-366) and FilmProgressive::put
//     (src/libaten/renderer/film.cpp:61-71);
//   * the shading / sampling math lives in aten_b200/csrc/atn_math.cuh (one text, compiled here
//     with g++ and by nvcc for the kernels); each function there cites its reference lines.
//
// Parity is PINNED: tests/test_oracle_cpu.py checks this restatement bit-exactly (hits, rays) and
// to 1e-6 (radiance) against vectors produced by the reference's own CPU renderer compiled in
// place (oracle/_ref, see oracle/build_ref.py and tests/golden/make_golden.py).
#include <omp.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "../aten_b200/csrc/atn_math.cuh"
#include "../aten_b200/csrc/atn_svgf.cuh"
#include "../aten_b200/csrc/atn_restir.cuh"

using namespace atn;

namespace {

int32_t g_device_texture_fetch = 0;     // oracle_set_texture_fetch: 1 = the GPU renderer's fetch semantics (atn_math.cuh tex_fetch_device)

struct Scene {
    SceneView view;
    std::vector<const AtnBvhNode*> lists;
    std::vector<TexView> tex;
    float t_min;
};

Scene MakeScene(const AtnSceneBlob* b)
{
    Scene s;
    s.view.objects = b->objects; s.view.matrices = b->matrices; s.view.materials = b->materials;
    s.view.lights = b->lights; s.view.triangles = b->triangles; s.view.positions = b->positions;
    s.view.normals = b->normals; s.view.light_count = (int32_t)b->light_count; s.view.material_count = (int32_t)b->material_count; s.view.cfg = b->config;
    for (uint32_t i = 0; i < b->texture_count; i++) {
        TexView t; t.texels = b->textures[i].texels; t.w = b->textures[i].width; t.h = b->textures[i].height;
        t.filter = b->textures[i].filter_mode; t.address = b->textures[i].address_mode;
        s.tex.push_back(t);
    }
    s.view.textures = s.tex.data();
    for (uint32_t i = 0; i < b->node_list_count; i++) s.lists.push_back(b->node_lists[i]);
    // ImageBasedLight::sample's scene radius (src/libaten/light/ibl.h:105-110, aabb.h:231-234,345-362)
    const float* mn = b->config.scene_bbox_min; const float* mx = b->config.scene_bbox_max;
    const bool valid = !((mn[0] >= mx[0]) || (mn[1] >= mx[1]) || (mn[2] >= mx[2]));
    float radius = 10000.0f;
    if (valid) {
        const V3 vmin(mn[0], mn[1], mn[2]), vmax(mx[0], mx[1], mx[2]);
        const V3 center = (vmin + vmax) * 0.5f;
        const float r = length(vmax - center);
        const float theta = kPi * 30.0f / 180.0f;
        radius = r / tanf(theta / 2);
    }
    s.view.scene_radius = radius;
    s.view.alpha_blending = b->enable_alpha_blending ? 1 : 0;
    s.view.device_texture_fetch = g_device_texture_fetch;
    s.t_min = b->config.bvh_hit_min > 0 ? b->config.bvh_hit_min : kEps;
    return s;
}

struct Counters { uint64_t nodes, boxes, tris, insts; };

// ThreadedBvhTraverser<true>::Traverse<Closest> (threaded_bvh_traverser.h:98-304); lod_depth = -1.
bool Traverse(const Scene& s, const Ray& r, float t_max, AtnHit* isect, Counters* cnt)
{
    const float t_min = s.t_min;
    isect->t = kInf; isect->objid = -1; isect->mtrlid = -1; isect->meshid = -1;
    isect->primid = -1; isect->a = 0; isect->b = 0; isect->isVoxel = 0;
    const AtnBvhNode* list = s.lists[0];
    Ray tr = r;
    RayInv ri = ray_inv(tr.org, tr.dir);
    int32_t nodeid = 0, objid = -1, meshid = -1, top_hit = -1, top_miss = -1;
    while (nodeid >= 0) {
        bool is_hit = false;
        AtnBvhNode node = list[nodeid];
        if (cnt) cnt->nodes++;
        if (node.object_id >= 0 || node.primid >= 0) {
            if (node.exid >= 0) {
                const AtnObjectParameter& obj = s.view.objects[(uint32_t)node.object_id];
                if (obj.mtx_id >= 0) {
                    const AtnMat4& W2L = s.view.matrices[obj.mtx_id + 1];
                    tr = make_ray(mat_apply(W2L, r.org), mat_apply_xyz(W2L, r.dir));   // mat4::applyRay
                }
                else tr = r;
                ri = ray_inv(tr.org, tr.dir);
                const uint32_t bits = (uint32_t)f2i(node.exid);
                const int32_t exid = (int32_t)(bits & 0x7fffu);
                list = s.lists[exid];
                objid = (int32_t)node.object_id;
                meshid = (int32_t)node.meshid;
                top_hit = (int32_t)node.hit;
                top_miss = (int32_t)node.miss;
                is_hit = true;
                node.hit = 0;
                if (cnt) cnt->insts++;
            }
            else if (node.primid >= 0) {
                const AtnTriangleParameter& prim = s.view.triangles[(uint32_t)node.primid];
                const V3 v0(s.view.positions[prim.idx[0]]), v1(s.view.positions[prim.idx[1]]), v2(s.view.positions[prim.idx[2]]);
                const TriHit h = intersect_triangle(tr.org, tr.dir, v0, v1, v2);
                if (cnt) cnt->tris++;
                is_hit = h.hit && h.t < kInf;          // triangle::hit against isect_tmp.t = INF
                const float tt = is_hit ? h.t : kInf;
                const bool is_intersect = t_min < tt && (is_hit && tt < isect->t);
                if (is_intersect) {
                    isect->t = h.t; isect->a = h.a; isect->b = h.b;
                    isect->objid = objid;
                    isect->primid = (int32_t)node.primid;
                    isect->mtrlid = prim.mtrlid;
                    isect->meshid = prim.mesh_id < 0 ? meshid : prim.mesh_id;
                    isect->isVoxel = 0;
                    t_max = isect->t;
                }
            }
        }
        else {
            float t0;
            is_hit = aabb_hit(ri, V3(node.boxmin[0], node.boxmin[1], node.boxmin[2]),
                              V3(node.boxmax[0], node.boxmax[1], node.boxmax[2]), t_min, t_max, &t0);
            if (cnt) cnt->boxes++;
        }
        nodeid = is_hit ? (int32_t)node.hit : (int32_t)node.miss;
        if (nodeid < 0) {
            nodeid = is_hit ? top_hit : top_miss;
            top_hit = -1; top_miss = -1;
            list = s.lists[0];
            tr = r;
            ri = ray_inv(tr.org, tr.dir);
        }
    }
    return isect->objid >= 0;
}

}  // namespace

extern "C" {

__attribute__((visibility("default")))
void oracle_set_texture_fetch(int32_t device_semantics) { g_device_texture_fetch = device_semantics ? 1 : 0; }

// tex_fetch_device on one texture for a list of coordinates (tests/test_texture_fetch.py pins it to an independent restatement)
__attribute__((visibility("default")))
void oracle_tex_fetch(const AtnVec4* texels, int32_t w, int32_t h, int32_t filter, int32_t address, const float* uv, int32_t n, AtnVec4* out)
{
    TexView t; t.texels = texels; t.w = w; t.h = h; t.filter = filter; t.address = address;
    for (int32_t i = 0; i < n; i++) out[i] = tex_fetch_device(t, uv[2 * i], uv[2 * i + 1]);
}

__attribute__((visibility("default")))
int oracle_trace_closest(const AtnSceneBlob* blob, const AtnRay* rays, const float* t_max, size_t n, AtnHit* hits,
                         uint64_t* counters /* [4] nodes, boxes, tris, insts; may be NULL */)
{
    const Scene s = MakeScene(blob);
    uint64_t cn = 0, cb = 0, ct = 0, ci = 0;
#pragma omp parallel for reduction(+:cn,cb,ct,ci) schedule(dynamic, 256)
    for (long i = 0; i < (long)n; i++) {
        Ray r; r.org = V3(rays[i].org); r.dir = V3(rays[i].dir);
        Counters c{0, 0, 0, 0};
        Traverse(s, r, t_max ? t_max[i] : kInf, &hits[i], counters ? &c : nullptr);
        cn += c.nodes; cb += c.boxes; ct += c.tris; ci += c.insts;
    }
    if (counters) { counters[0] = cn; counters[1] = cb; counters[2] = ct; counters[3] = ci; }
    return 0;
}

// Shadow query with the semantics of atn_trace_any (singular light at distance dist[i]).
__attribute__((visibility("default")))
int oracle_trace_any(const AtnSceneBlob* blob, const AtnRay* rays, const float* dist, size_t n, uint8_t* visible)
{
    const Scene s = MakeScene(blob);
#pragma omp parallel for schedule(dynamic, 256)
    for (long i = 0; i < (long)n; i++) {
        Ray r; r.org = V3(rays[i].org); r.dir = V3(rays[i].dir);
        AtnHit h;
        const bool hit = Traverse(s, r, dist[i] - kEps, &h, nullptr);
        visible[i] = (!hit || h.t > dist[i]) ? 1 : 0;
    }
    return 0;
}

// aten::PathTracing::OnRender driven as `frames` successive frames of one sample, with
// FilmProgressive accumulation; `records` (optional) receives AtnBounceRecord[pixel][max_depth]
// of the LAST frame; stats[6] = {closest rays, shadow rays, nodes, slab tests, triangle tests, instance entries}.
__attribute__((visibility("default")))
int oracle_render(const AtnSceneBlob* blob, const AtnCameraParameter* cam, const uint32_t* seeds, uint32_t seed_count,
                  int32_t W, int32_t H, uint32_t first_frame, int32_t frames, int32_t max_depth, int32_t rr_depth,
                  AtnVec4* image /* in/out progressive film, W*H */, AtnBounceRecord* records, uint64_t* stats,
                  AtnVec4* aov_normal_depth /* W*H or NULL */, AtnVec4* aov_albedo_meshid /* W*H or NULL */)
{
    const Scene s = MakeScene(blob);
    if (rr_depth > max_depth) rr_depth = max_depth - 1;
    uint64_t n_closest = 0, n_shadow = 0, cn = 0, cb = 0, ct = 0, ci = 0;
    const bool want_cnt = stats != nullptr;
    for (int32_t f = 0; f < frames; f++) {
        const uint32_t frame = first_frame + (uint32_t)f;
        AtnBounceRecord* recs = (f == frames - 1) ? records : nullptr;
#pragma omp parallel for reduction(+:n_closest,n_shadow,cn,cb,ct,ci) schedule(dynamic, 4)
        for (int32_t y = 0; y < H; y++) {
            Counters cnt{0, 0, 0, 0};
            for (int32_t x = 0; x < W; x++) {
                const int32_t idx = y * W + x;
                const uint32_t rnd = seeds[(uint32_t)idx % seed_count];
                PathState st;
                st.sampler = sampler_for_path(frame, rnd);
                const float r1 = st.sampler.next();
                const float r2 = st.sampler.next();
                const float sx = (x + r1) / (float)(cam->width);
                const float sy = (y + r2) / (float)(cam->height);
                Ray ray = camera_sample(*cam, sx, sy);
                st.throughput = V3(1.0f); st.pdfb = 1.0f; st.contrib = V3(0.0f);
                st.terminated = false; st.singular = false;
                int32_t depth = 0;
                while (depth < max_depth) {
                    AtnHit isect;
                    const bool is_hit = Traverse(s, ray, kInf, &isect, want_cnt ? &cnt : nullptr);
                    n_closest++;
                    AtnBounceRecord rec;
                    memset(&rec, 0, sizeof(rec));
                    rec.ray.org = AtnVec3{ray.org.x, ray.org.y, ray.org.z};
                    rec.ray.dir = AtnVec3{ray.dir.x, ray.dir.y, ray.dir.z};
                    rec.hit = isect;
                    rec.flags = 1;
                    bool will_continue = true;
                    if (is_hit) {
                        ShadowRayOut sh;
                        Ray next = ray;
                        // shade_hit() == the three stages in a row; the first one also stores the AOVs of the
                        // first hit the way the reference's GPU kernel does (pathtracing_impl.cu:92-111)
                        sh.active = false;
                        Surface sf;
                        const bool lit = shade_surface(s.view, ray, isect.objid, isect.primid, isect.a, isect.b, isect.t, isect.mtrlid,
                                                       depth, &st, &sf,
                                                       aov_normal_depth ? &aov_normal_depth[idx] : nullptr,
                                                       aov_albedo_meshid ? &aov_albedo_meshid[idx] : nullptr, isect.meshid);
                        if (lit) {
                            shade_nee(s.view, ray.dir, sf, &st, &sh);
                            shade_bsdf(s.view, ray.dir, sf, depth, rr_depth, &st, &next);
                        }
                        // HitShadowRay (pathtracing_impl.h:353-393)
                        if (!st.terminated && sh.active) {
                            n_shadow++;
                            Ray sr; sr.org = sh.org; sr.dir = normalize(sh.dir);   // aten::ray(org, dir) renormalises
                            // HitTestToTargetLight's loop (pathtracing_impl.h:288-348): one lookup unless alpha blending is on
                            // or the surface has stencil type ALWAYS
                            const AtnLightParameter& light = s.view.lights[sh.light_id];
                            const bool need_stencil = shadow_needs_stencil(s.view, isect.mtrlid);
                            const int lookups = shadow_max_lookups(s.view, need_stencil);
                            const V3 orig_dir = sr.dir;
                            int32_t hitobj = shadow_light_obj(light);
                            bool vis = false;
                            for (int k = 0; k < lookups; k++) {
                                AtnHit shit;
                                Traverse(s, sr, sh.dist - kEps, &shit, want_cnt ? &cnt : nullptr);
                                V3 next_org;
                                const int res = shadow_lookup(s.view, light, sr, orig_dir, sh.dist, shit.objid, shit.primid, shit.a, shit.b,
                                                              shit.t, shit.mtrlid, k + 1 >= lookups, need_stencil, &hitobj, &next_org);
                                if (res != kShadowRetrace) { vis = res == kShadowVisible; break; }
                                sr.org = next_org; sr.dir = normalize(orig_dir);      // aten::ray(rec.p, original_ray.dir, normal)
                            }
                            if (vis) st.contrib = st.contrib + sh.lightcontrib;
                            rec.shadow_active = 1;
                            rec.shadow_org[0] = sh.org.x; rec.shadow_org[1] = sh.org.y; rec.shadow_org[2] = sh.org.z;
                            rec.shadow_dist = sh.dist;
                            rec.shadow_dir[0] = sh.dir.x; rec.shadow_dir[1] = sh.dir.y; rec.shadow_dir[2] = sh.dir.z;
                            rec.light_contrib[0] = sh.lightcontrib.x; rec.light_contrib[1] = sh.lightcontrib.y; rec.light_contrib[2] = sh.lightcontrib.z;
                            rec.shadow_visible = vis ? 1 : 0;
                        }
                        will_continue = !st.terminated;
                        ray = next;
                    }
                    else {
                        shade_miss(s.view, *cam, x, y, W, H, depth, ray, &st,
                                   aov_normal_depth ? &aov_normal_depth[idx] : nullptr,
                                   aov_albedo_meshid ? &aov_albedo_meshid[idx] : nullptr);
                        will_continue = false;
                    }
                    rec.throughput[0] = st.throughput.x; rec.throughput[1] = st.throughput.y; rec.throughput[2] = st.throughput.z;
                    rec.pdfb = st.pdfb;
                    rec.contrib[0] = st.contrib.x; rec.contrib[1] = st.contrib.y; rec.contrib[2] = st.contrib.z;
                    if (!will_continue) rec.flags |= 2;
                    if (recs) recs[(size_t)idx * max_depth + depth] = rec;
                    if (!will_continue) break;
                    depth++;
                }
                // dst.buffer->put(x, y, vec4(col / cnt, 1)) with cnt = 1, FilmProgressive::put
                AtnVec4& cur = image[idx];
                const float n = (float)((int32_t)cur.w);
                const V3 col = st.contrib / 1.0f;
                cur.x = (n * cur.x + col.x) / (n + 1);
                cur.y = (n * cur.y + col.y) / (n + 1);
                cur.z = (n * cur.z + col.z) / (n + 1);
                cur.w = n + 1;
            }
            cn += cnt.nodes; cb += cnt.boxes; ct += cnt.tris; ci += cnt.insts;
        }
    }
    // stats: closest rays, shadow rays, then the traversal counters of ALL rays on the reference
    // layout (nodes fetched, slab tests, triangle tests, instance entries) -- the roofline denominator
    if (stats) { stats[0] = n_closest; stats[1] = n_shadow; stats[2] = cn; stats[3] = cb; stats[4] = ct; stats[5] = ci; }
    return 0;
}

// aten::AORenderer::RenderAO (src/libaten/renderer/ao/aorenderer.cpp:75-146) driven as `frames` frames
// of one sample with FilmProgressive accumulation.
// host_row_break = 1 reproduces the stock host renderer bit for bit, including its quirk: the
// `if (is_terminated) break;` at aorenderer.cpp:136-138 sits in the x loop (there is no sample loop), so
// the first primary-ray miss of a scanline ends that scanline and neither that pixel nor the rest of the
// row is written to the film.  host_row_break = 0 is the per-pixel semantics of the reference's GPU
// kernels (src/libidaten/ao/ao.cu:20-35): a miss stores ShadeByAOIfHitMiss() = 1.0 and every pixel is
// independent -- what the CUDA path implements.
__attribute__((visibility("default")))
int oracle_render_ao(const AtnSceneBlob* blob, const AtnCameraParameter* cam, const uint32_t* seeds, uint32_t seed_count,
                     int32_t W, int32_t H, uint32_t first_frame, int32_t frames, int32_t num_rays, float radius,
                     int32_t host_row_break, AtnVec4* image)
{
    const Scene s = MakeScene(blob);
    for (int32_t f = 0; f < frames; f++) {
        const uint32_t frame = first_frame + (uint32_t)f;
#pragma omp parallel for schedule(dynamic, 4)
        for (int32_t y = 0; y < H; y++) {
            for (int32_t x = 0; x < W; x++) {
                const int32_t idx = y * W + x;
                const uint32_t rnd = seeds[(uint32_t)idx % seed_count];
                Sampler sp = sampler_for_path(frame, rnd);
                const float r1 = sp.next();
                const float r2 = sp.next();
                const Ray ray = camera_sample(*cam, (x + r1) / (float)(cam->width), (y + r2) / (float)(cam->height));
                sp = ao_sampler_for_path(frame, rnd);
                AtnHit isect;
                float ao = 1.0f;
                const bool is_hit = Traverse(s, ray, kInf, &isect, nullptr);
                if (!is_hit && host_row_break) break;
                if (is_hit) {
                    ao = 0.0f;
                    V3 p, n;
                    ao_surface(s.view, ray, isect.objid, isect.primid, isect.a, isect.b, isect.t, isect.mtrlid, &p, &n);
                    for (int32_t i = 0; i < num_rays; i++) {
                        V3 next_dir;
                        Ray ar = ao_ray(p, n, &sp, &next_dir);
                        for (int k = 0; k < 10; k++) {          // MAX_LOOP (aorenderer_impl.h:68)
                            AtnHit ah;
                            Traverse(s, ar, radius, &ah, nullptr);
                            Ray next;
                            if (!ao_accumulate(s.view, &ao, n, next_dir, ar, radius, ah.objid, ah.primid, ah.a, ah.b, ah.t, ah.mtrlid, k == 9, &next)) break;
                            ar = next;
                        }
                    }
                    ao /= num_rays;
                }
                AtnVec4& cur = image[idx];
                const float nn = (float)((int32_t)cur.w);
                const V3 col = V3(ao) / 1.0f;
                cur.x = (nn * cur.x + col.x) / (nn + 1);
                cur.y = (nn * cur.y + col.y) / (nn + 1);
                cur.z = (nn * cur.z + col.z) / (nn + 1);
                cur.w = nn + 1;
            }
        }
    }
    return 0;
}

__attribute__((visibility("default")))
// N1: aten::SVGFRenderer::OnRender (src/libaten/renderer/svgf/svgf.cpp:461-639), `frames` frames of one
// sample each, starting at frame 0 (frame 0 has no history).  State lives in `state`, 11 * W*H AtnVec4 laid
// out as: aov[0] {normal_depth, albedo_meshid, color_variance, moment_tw}, aov[1] {...}, atrous ping, atrous pong,
// temporary colour; the caller passes it zero-filled... except that the reference's buffers start as default
// aten::vec4 = (0,0,0,1), which this function writes itself when first_frame == 0.
// `motion_depth`: W*H motion vectors (xy, in screen fractions) or NULL for a static camera.
// `host_order` = 1: variance estimation updates the colour history in place in row-major order on one thread,
// which is what the stock host renderer does with one OpenMP thread; 0: order-independent form (all pixels
// read the unfiltered colours), which is what the GPU path computes.
// `image`: W*H film colours of the LAST frame rendered (last a-trous iteration, albedo put back).
int oracle_render_svgf(const AtnSceneBlob* blob, const AtnCameraParameter* cam, const uint32_t* seeds, uint32_t seed_count,
                       int32_t W, int32_t H, uint32_t first_frame, int32_t frames, int32_t max_depth, int32_t rr_depth,
                       const AtnVec4* motion_depth, int32_t host_order, int32_t atrous_iters, AtnVec4* state, AtnVec4* image)
{
    const Scene s = MakeScene(blob);
    if (rr_depth > max_depth) rr_depth = max_depth - 1;
    const size_t n = (size_t)W * H;
    SvgfAov aovs[2];
    for (int k = 0; k < 2; k++) {
        aovs[k].normal_depth = state + (k * 4 + 0) * n; aovs[k].albedo_meshid = state + (k * 4 + 1) * n;
        aovs[k].color_variance = state + (k * 4 + 2) * n; aovs[k].moment_tw = state + (k * 4 + 3) * n;
    }
    AtnVec4* atrous[2] = { state + 8 * n, state + 9 * n };
    AtnVec4* temporary = state + 10 * n;
    if (first_frame == 0)
        for (size_t i = 0; i < 11 * n; i++) { state[i].x = 0; state[i].y = 0; state[i].z = 0; state[i].w = 1; }
    std::vector<AtnVec4> contribs(n), scratch(host_order ? 0 : n);
    const AtnMat4 w2c = camera_w2c(*cam);
    const float camera_distance = screen_distance(*cam, H);
    const int32_t iter_count = atrous_iters > 0 ? atrous_iters : 5;      // SVGFParams::atrous_iter_cnt, svgf_types.h:74

    for (int32_t f = 0; f < frames; f++) {
        const uint32_t frame = first_frame + (uint32_t)f;
        SvgfAov cur = aovs[frame & 1], prev = aovs[1 - (frame & 1)];
        // ---- one sample of path tracing per pixel (ExecRendering + Shade) + PrepareForDenoise ----
#pragma omp parallel for schedule(dynamic, 4)
        for (int32_t y = 0; y < H; y++) {
            for (int32_t x = 0; x < W; x++) {
                const int32_t idx = y * W + x;
                const uint32_t rnd = seeds[(uint32_t)idx % seed_count];
                PathState st;
                st.sampler = sampler_for_path(frame, rnd);
                const float r1 = st.sampler.next();
                const float r2 = st.sampler.next();
                Ray ray = camera_sample(*cam, (x + r1) / (float)(cam->width), (y + r2) / (float)(cam->height));
                st.throughput = V3(1.0f); st.pdfb = 1.0f; st.contrib = V3(0.0f);
                st.terminated = false; st.singular = false; st.last_specular = false;
                for (int32_t depth = 0; depth < max_depth; depth++) {
                    AtnHit isect;
                    if (!Traverse(s, ray, kInf, &isect, nullptr)) {
                        shade_miss(s.view, *cam, x, y, W, H, depth, ray, &st, &cur.normal_depth[idx], &cur.albedo_meshid[idx]);
                        break;
                    }
                    ShadowRayOut sh;
                    Ray next = ray;
                    shade_hit(s.view, ray, isect.objid, isect.primid, isect.a, isect.b, isect.t, isect.mtrlid, depth, rr_depth,
                              &st, &sh, &next, &cur.normal_depth[idx], &cur.albedo_meshid[idx], &w2c);
                    if (!st.terminated && sh.active) {
                        Ray sr; sr.org = sh.org; sr.dir = normalize(sh.dir);
                        AtnHit shit;
                        Traverse(s, sr, sh.dist - kEps, &shit, nullptr);
                        if (shadow_visible(s.view, s.view.lights[sh.light_id], sr, sh.dist, shit.objid, shit.primid, shit.a, shit.b, shit.t, shit.mtrlid))
                            st.contrib = st.contrib + sh.lightcontrib;
                    }
                    if (st.terminated) break;
                    ray = next;
                }
                AtnVec4 c; c.x = st.contrib.x; c.y = st.contrib.y; c.z = st.contrib.z; c.w = 1.0f;   // one sample
                contribs[idx] = c;
                svgf_prepare(frame == 0, idx, c, temporary, cur);
            }
        }
        // ---- temporal reprojection ----
        if (frame > 0) {
#pragma omp parallel for schedule(static)
            for (int32_t y = 0; y < H; y++)
                for (int32_t x = 0; x < W; x++)
                    svgf_temporal(x, y, W, H, 0.98f, 0.05f, contribs[y * W + x], cur, prev, motion_depth);
        }
        // ---- variance estimation ----
        if (host_order) {
            for (int32_t y = 0; y < H; y++)
                for (int32_t x = 0; x < W; x++)
                    svgf_variance(x, y, W, H, camera_distance, cur, cur.color_variance, cur.color_variance);
        }
        else {
#pragma omp parallel for schedule(static)
            for (int32_t y = 0; y < H; y++)
                for (int32_t x = 0; x < W; x++)
                    svgf_variance(x, y, W, H, camera_distance, cur, cur.color_variance, scratch.data());
            memcpy(cur.color_variance, scratch.data(), n * sizeof(AtnVec4));
        }
        // ---- a-trous iterations ----
        for (int32_t i = 0; i < iter_count; i++) {
            const AtnVec4* src = (i == 0) ? cur.color_variance : atrous[i & 1];
            AtnVec4* next = atrous[1 - (i & 1)];
#pragma omp parallel for schedule(static)
            for (int32_t y = 0; y < H; y++)
                for (int32_t x = 0; x < W; x++) {
                    AtnVec4 film;
                    if (svgf_atrous(i, iter_count, x, y, W, H, camera_distance, cur, src, next, temporary, &film)) image[y * W + x] = film;
                }
        }
        for (size_t i = 0; i < n; i++) svgf_copy((int32_t)i, temporary, cur);
    }
    return 0;
}

// N1: the ReSTIR renderer -- the reference's per-pixel steps (restir_impl.h, ReSTIRRenderer::Shade) in the order of
// idaten::ReSTIRPathTracing::OnRender, frames 0 .. frames-1 of one sample each (oracle/ref_harness.cpp, ReplayReSTIRFrame,
// is the same sequence built from the reference's own functions).  `motion_depth`: W*H motion vectors or NULL.
// `image`: W*H contribution of the LAST frame (rgb, 1).  `reservoirs_out`: W*H x 5 floats (w_sum, M, y, W, target pdf)
// of the reservoirs that frame's pixel colours were computed from, or NULL.
int oracle_render_restir(const AtnSceneBlob* blob, const AtnCameraParameter* cam, const uint32_t* seeds, uint32_t seed_count,
                         int32_t W, int32_t H, int32_t frames, int32_t max_depth, int32_t rr_depth,
                         const AtnVec4* motion_depth, AtnVec4* image, float* reservoirs_out)
{
    const Scene s = MakeScene(blob);
    if (rr_depth > max_depth) rr_depth = max_depth - 1;
    const int32_t n = W * H;
    std::vector<Reservoir> res[2] = { std::vector<Reservoir>(n), std::vector<Reservoir>(n) };
    std::vector<RestirInfo> infos[2] = { std::vector<RestirInfo>(n), std::vector<RestirInfo>(n) };
    for (int k = 0; k < 2; k++)
        for (int32_t i = 0; i < n; i++) {
            memset((void*)&res[k][i], 0, sizeof(Reservoir)); res[k][i].y = -1;               // Reservoir's default member initialisers
            memset((void*)&infos[k][i], 0, sizeof(RestirInfo)); infos[k][i].mtrl_idx = -1; infos[k][i].mesh_id = -1;
        }
    std::vector<V3> stale_lightcontrib(n, V3(0.0f));      // ShadowRay::lightcontrib, kept across bounces and frames
    std::vector<AtnVec4> aov_nd(n), aov_am(n);
    std::vector<PathState> paths(n);
    std::vector<Ray> rays(n);
    const AtnMat4 w2c = camera_w2c(*cam);
    int cur = 0, target = 0;

    for (int32_t frame = 0; frame < frames; frame++) {
        for (int32_t idx = 0; idx < n; idx++) {
            const int32_t x = idx % W, y = idx / W;
            PathState& st = paths[idx];
            st.sampler = sampler_for_path((uint32_t)frame, seeds[(uint32_t)idx % seed_count]);
            const float r1 = st.sampler.next();
            const float r2 = st.sampler.next();
            rays[idx] = camera_sample(*cam, (x + r1) / (float)(cam->width), (y + r2) / (float)(cam->height));
            st.throughput = V3(1.0f); st.pdfb = 1.0f; st.contrib = V3(0.0f);
            st.terminated = false; st.singular = false; st.last_specular = false;
            res[cur][idx].clear();
            infos[cur][idx].clear();
        }
        for (int32_t bounce = 0; bounce < max_depth; bounce++) {
#pragma omp parallel for schedule(dynamic, 64)
            for (int32_t idx = 0; idx < n; idx++) {
                PathState& st = paths[idx];
                if (st.terminated) continue;
                const Ray ray = rays[idx];
                AtnHit isect;
                if (!Traverse(s, ray, kInf, &isect, nullptr)) {
                    shade_miss(s.view, *cam, idx % W, idx / W, W, H, bounce, ray, &st, &aov_nd[idx], &aov_am[idx]);
                    continue;
                }
                if (bounce == 0) {
                    Surface sf;
                    if (!restir_surface(s.view, ray, isect.objid, isect.primid, isect.a, isect.b, isect.t, isect.mtrlid, isect.meshid,
                                        bounce, w2c, &st, &sf, &infos[cur][idx], &aov_nd[idx], &aov_am[idx]))
                        continue;
                    const Mtrl mtrl = fill_material(s.view, sf.mtrlid);
                    if (!(mtrl.is_singular() || mtrl.is_translucent()))
                        restir_initial_candidate(&res[cur][idx], mtrl, s.view, sf.p, sf.n, ray.dir, sf.u, sf.v, &st.sampler);
                    Ray next = ray;
                    shade_bsdf(s.view, ray.dir, sf, bounce, rr_depth, &st, &next);
                    rays[idx] = next;
                }
                else {
                    ShadowRayOut sh;
                    Ray next = ray;
                    shade_hit(s.view, ray, isect.objid, isect.primid, isect.a, isect.b, isect.t, isect.mtrlid, bounce, rr_depth, &st, &sh, &next);
                    if (sh.sampled) stale_lightcontrib[idx] = sh.lightcontrib;
                    if (!st.terminated && sh.active) {
                        Ray sr; sr.org = sh.org; sr.dir = normalize(sh.dir);
                        AtnHit shit;
                        Traverse(s, sr, sh.dist - kEps, &shit, nullptr);
                        if (shadow_visible(s.view, s.view.lights[sh.light_id], sr, sh.dist, shit.objid, shit.primid, shit.a, shit.b, shit.t, shit.mtrlid))
                            st.contrib = st.contrib + sh.lightcontrib;
                    }
                    rays[idx] = next;
                }
            }
            if (bounce > 0) continue;
            // visibility of every reservoir's sample
#pragma omp parallel for schedule(dynamic, 64)
            for (int32_t idx = 0; idx < n; idx++) {
                PathState& st = paths[idx];
                if (st.terminated) continue;
                Reservoir& r = res[cur][idx];
                bool reached = false;
                if (r.valid()) {
                    V3 org, dir; float dist;
                    restir_shadow_ray(r, &infos[cur][idx], &org, &dir, &dist);
                    Ray sr; sr.org = org; sr.dir = normalize(dir);       // aten::ray(org, dir) renormalises
                    AtnHit shit;
                    Traverse(s, sr, dist - kEps, &shit, nullptr);
                    reached = shadow_visible(s.view, s.view.lights[r.y], sr, dist, shit.objid, shit.primid, shit.a, shit.b, shit.t, shit.mtrlid);
                    if (reached) st.contrib = st.contrib + stale_lightcontrib[idx];
                }
                restir_visibility_result(&r, reached);
            }
            const int dst = (cur + 1) & 1, prev = (cur + 1) & 1;
            const int cur_infos = cur;
            if (frame > 1) {
#pragma omp parallel for schedule(dynamic, 64)
                for (int32_t idx = 0; idx < n; idx++) {
                    if (paths[idx].terminated) continue;
                    restir_temporal(idx, W, H, s.view, &paths[idx].sampler, &res[cur][idx], infos[cur][idx], res[prev].data(), infos[prev].data(),
                                    aov_am.data(), motion_depth);
                }
            }
#pragma omp parallel for schedule(dynamic, 64)
            for (int32_t idx = 0; idx < n; idx++) {
                if (paths[idx].terminated) continue;
                restir_spatial(idx, W, H, s.view, &paths[idx].sampler, &res[dst][idx], res[cur].data(), infos[cur].data(), aov_am.data());
            }
            for (int32_t idx = 0; idx < n; idx++) {
                if (paths[idx].terminated) continue;
                V3 c;
                if (restir_pixel_color(s.view, res[dst][idx], infos[cur_infos][idx], aov_am[idx], &c)) paths[idx].contrib = paths[idx].contrib + c;
            }
            target = dst;
            cur = (cur + 1) & 1;      // ReuseParams::Update
        }
        for (int32_t idx = 0; idx < n; idx++) {
            const V3 c = paths[idx].contrib;
            image[idx].x = c.x; image[idx].y = c.y; image[idx].z = c.z; image[idx].w = 1.0f;
        }
    }
    if (reservoirs_out)
        for (int32_t idx = 0; idx < n; idx++) {
            const Reservoir& r = res[target][idx];
            float* o = reservoirs_out + (size_t)idx * 5;
            o[0] = r.w_sum; o[1] = (float)r.M; o[2] = (float)r.y; o[3] = r.W; o[4] = r.target_pdf;
        }
    return 0;
}

int oracle_cmj(uint32_t idx, uint32_t scramble, int32_t n, float* out)
{
    Sampler s; s.idx = idx; s.dim = 0; s.scramble = scramble;
    for (int32_t i = 0; i < n; i++) out[i] = s.next();
    return 0;
}

// Camera rays of GeneratePath (pathtracing_impl.h:65-110) for a whole frame.
__attribute__((visibility("default")))
int oracle_camera_rays(const AtnCameraParameter* cam, const uint32_t* seeds, uint32_t seed_count,
                       int32_t W, int32_t H, uint32_t frame, AtnRay* rays)
{
    for (int32_t y = 0; y < H; y++) for (int32_t x = 0; x < W; x++) {
        const int32_t idx = y * W + x;
        Sampler sp = sampler_for_path(frame, seeds[(uint32_t)idx % seed_count]);
        const float r1 = sp.next(), r2 = sp.next();
        const Ray r = camera_sample(*cam, (x + r1) / (float)cam->width, (y + r2) / (float)cam->height);
        rays[idx].org = AtnVec3{r.org.x, r.org.y, r.org.z};
        rays[idx].dir = AtnVec3{r.dir.x, r.dir.y, r.dir.z};
    }
    return 0;
}

// The reference's own known-answer tests for this path, restated as callable functions:
// FillBasicAOVs / FillBasicAOVsIfHitMiss (src/aten_unittest/aov_host_buffer.cpp:73-130) and
// Camera::ComputePixelWidthAtDistance (src/aten_unittest/pinhole_camera.cpp:6-15).
__attribute__((visibility("default")))
void oracle_fill_basic_aovs(const float normal[3], const float p[3], const float w2c[16], const float albedo[4], int32_t meshid,
                            float normal_depth[4], float albedo_meshid[4])
{
    AtnMat4 m; memcpy(m.m, w2c, sizeof(m.m));
    AtnVec4 al{ albedo[0], albedo[1], albedo[2], albedo[3] };
    AtnVec4 nd, am;
    fill_basic_aovs(&nd, V3(normal[0], normal[1], normal[2]), V3(p[0], p[1], p[2]), m, &am, al, meshid);
    memcpy(normal_depth, &nd, 16); memcpy(albedo_meshid, &am, 16);
}

__attribute__((visibility("default")))
void oracle_fill_basic_aovs_miss(const float bg[4], float normal_depth[4], float albedo_meshid[4])
{
    AtnVec4 b{ bg[0], bg[1], bg[2], bg[3] }, nd, am;
    fill_basic_aovs_miss(&nd, &am, b);
    memcpy(normal_depth, &nd, 16); memcpy(albedo_meshid, &am, 16);
}

__attribute__((visibility("default")))
float oracle_pixel_width_at_distance(float vfov, int32_t width, int32_t height, float distance)
{
    AtnCameraParameter c; memset(&c, 0, sizeof(c));
    c.vfov = vfov; c.width = width; c.height = height;
    return pixel_width_at_distance(c, distance);
}

__attribute__((visibility("default")))
void oracle_set_threads(int n) { omp_set_num_threads(n); }

}  // extern "C"
