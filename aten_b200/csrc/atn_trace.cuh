// atn_trace.cuh -- the traversal kernels of atn_kernels.cu (one translation unit; included once, after the constants,
// Counters and PathBuffers): ray / hit policies, k_trace<Policy> (persistent warp-synchronous traversal over QuadNodes, default
// and streaming form), and the exact walk of near-tie rays over the reference's own node lists (ref_traverse, k_exact).
#pragma once

// ---------------------------------------------------------------------------------------------
// traversal
// ---------------------------------------------------------------------------------------------
struct TraceResult {
    float t, a, b;
    int32_t leaf;      // tri leaf index (tie-break rank), -1 none
    int32_t inst;      // instance leaf index
};

enum : int { EXIT_NONE = 0, EXIT_ANY_HIT = 1, EXIT_HIT_BEFORE = 2 };

__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

// Resolve the winning leaf into the reference's Intersection fields.
__device__ __forceinline__ void resolve_hit(const BvhView& bv, const TraceResult& r, int32_t* objid, int32_t* primid,
                                            int32_t* mtrlid, int32_t* meshid)
{
    if (r.leaf < 0) { *objid = -1; *primid = -1; *mtrlid = -1; *meshid = -1; return; }
    const float4* tp = bv.tris + (size_t)r.leaf * kTriStride;
    *primid = __float_as_int(ldg4(tp + 0).w);
    *mtrlid = __float_as_int(ldg4(tp + 1).w);
    const int32_t mesh = __float_as_int(ldg4(tp + 2).w);
    const int4 il = __ldg(reinterpret_cast<const int4*>(bv.insts) + (size_t)r.inst * kInstStride);
    *objid = il.x;
    *meshid = mesh < 0 ? il.w : mesh;
}

// ---- near ties ---------------------------------------------------------------------------------------------
// The reference culls boxes with its RUNNING closest distance through a slab test that is biased by +1e-6 on the
// direction (aabb.h:68) and walks its trees in a fixed depth-first order (threaded_bvh_traverser.h:98-304).  A box
// can therefore appear to start BEHIND a triangle it contains, and when two primitives are hit within that error of
// each other, which of them the reference reports depends on its visit order: it may keep the farther one because
// the nearer one's box was culled by then.  This kernel walks front to back, so for such rays -- and only for them --
// its own answer can be the other primitive.  They are detected here and decided by the reference's own walk:
//   * per ray a band  B(t) = |t| * rho + alpha  bounds how far the biased entry distance of a box can lie beyond the
//     computed hit distance of a triangle inside it:
//         rho   = max(kBandRel, kBandInv * 1e-6 * max_k |invdir_k|)      the bias 1e-6 / |dir_k + 1e-6| and rounding,
//         alpha = kBandS * sum_k |oxinvdir_k| + kBandM * (1 + max_k |org_k|)   rounding of plane * inv - org * inv and of
//                                                                         the triangle test, absolute,
//     taken over the world ray and every instance-space ray entered so far;
//   * boxes are culled with t + 2 B(t), so every primitive that can tie with the final hit has been tested;
//   * the smallest hit distance of a primitive OTHER than the best one is tracked (t2); at the end of the ray, t2 within
//     B of the best hit means the reference's answer depends on its order: the ray goes to the exact queue and
//     k_exact repeats it as the reference's own depth-first walk over the reference's own node lists.
//   * most ties are still decided here: stack keys carry the LARGEST entry distance on the path from the root (pathmax), and
//     if the best hit is strictly nearer than every runner-up (t_best < t2) and every box on the path to its leaf starts
//     no later than t2 (pathmax <= t2), the reference reaches that leaf whatever it has accepted before -- everything it
//     can have accepted is at t2 or beyond -- and keeps it.  Only exact ties (t_best == t2) and paths with a box that
//     starts behind t2 go to the exact walk.
// If nothing ties, the reference's answer is this kernel's: every other primitive is farther than the band, the boxes on
// the path to the best primitive start no later than its hit plus the band, so no hit the reference accepts earlier
// can cull them (DESIGN.md "Exactness").  scripts/exp_near_ties.py measures the rule on the CPU emulation: with these
// constants no ray that differs from the reference goes unflagged (Sponza: 0.4 % of path rays flagged, 0.01 % differ).
constexpr float kBandRel = 5e-5f, kBandInv = 3.0f, kBandS = 2e-6f, kBandM = 5e-5f;
struct Band { float rho, alpha; };
__device__ __forceinline__ Band band_of(const V3& org, const RayInv& ri)
{
    const float im = fmaxf(fmaxf(fabsf(ri.invdir.x), fabsf(ri.invdir.y)), fabsf(ri.invdir.z));
    const float S = fabsf(ri.oxinvdir.x) + fabsf(ri.oxinvdir.y) + fabsf(ri.oxinvdir.z);
    const float M = fmaxf(fmaxf(fabsf(org.x), fabsf(org.y)), fabsf(org.z));
    Band b;
    b.rho = fmaxf(kBandRel, kBandInv * 1e-6f * im);
    b.alpha = kBandS * S + kBandM * (1.0f + M);
    return b;
}
__device__ __forceinline__ float band_at(const Band& b, float t) { return fabsf(t) * b.rho + b.alpha; }
// bound used to CULL boxes once a hit at distance t is known
__device__ __forceinline__ float cull_bound(const Band& b, float t) { return fminf(t + 2.0f * band_at(b, t), kInf); }

// What the traversal knows about runners-up when a ray ends.
struct Tie {
    bool any;      // another primitive was hit within `band` of the best hit
    bool cls;      // ... one of another instance or another material (only tracked for Policy::kNeedClass)
    float band;    // B(best.t)
};

// intersectTriangle (src/libaten/math/intersect.h:45-89) with e1 = v1 - v0 and e2 = v2 - v0 taken from
// the TriLeaf record (computed on upload with the same float subtraction, so every later operation
// sees the same operands as the reference's).
__device__ __forceinline__ TriHit intersect_triangle_edges(const V3& org, const V3& dir, const V3& v0, const V3& e1, const V3& e2)
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

// aabb::hit (src/libaten/math/aabb.h:61-86) for the kernels: same products, sums and comparisons as
// atn::aabb_hit, with the min/max selections as single FMNMX instructions.  fminf/fmaxf differ from
// the reference's std::min/std::max only when an operand is NaN, which needs a direction component
// equal to -1e-6f exactly (1/(dir + 1e-6) = inf); DESIGN.md lists this as not reproduced.
__device__ __forceinline__ bool slab_test(const RayInv& ri, float minx, float maxx, float miny, float maxy,
                                          float minz, float maxz, float t_min, float t_max, float* t0_out, float* t1_out)
{
    const float fx = maxx * ri.invdir.x + ri.oxinvdir.x;
    const float fy = maxy * ri.invdir.y + ri.oxinvdir.y;
    const float fz = maxz * ri.invdir.z + ri.oxinvdir.z;
    const float nx = minx * ri.invdir.x + ri.oxinvdir.x;
    const float ny = miny * ri.invdir.y + ri.oxinvdir.y;
    const float nz = minz * ri.invdir.z + ri.oxinvdir.z;
    const float t1raw = fminf(fminf(fmaxf(fx, nx), fmaxf(fy, ny)), fmaxf(fz, nz));
    const float t0 = fmaxf(fmaxf(fmaxf(fminf(fx, nx), fminf(fy, ny)), fminf(fz, nz)), t_min);
    *t0_out = t0;
    *t1_out = t1raw;                       // exit distance, not clamped by t_max (leaf filter uses its own bound)
    return t0 <= fminf(t1raw, t_max);
}

// Instance pre-filter.  The reference enters every instance whose parent box in the top layer was hit
// (leaves are never box-tested, threaded_bvh_traverser.h:146-185); the entry costs a matrix transform
// and a root test that run for one or two lanes at a time.  InstLeafW carries a world-space box that is
// conservative BY CONSTRUCTION (the instance's transformed root box, padded, atn_bvh.h), and the kernel
// skips the entry when the world ray misses that box by a safe margin: whatever the reference would
// accept inside the instance lies (to float accuracy) inside that box and passes the relaxed test.
// Far bound: the running hit distance only (the reference's triangle test ignores the caller's t_max,
// threaded_bvh_traverser.h:198-200).
//   margin in t:  2e-4 (|t0| + |t1|)          covers the +1e-6 direction bias, <= 5e-5 relative for |d_k| >= 0.02
//               + 1e-6 sum_k |org_k / d_k|    covers the rounding of plane*inv - org*inv
//               + 5e-4 (1 + max_k |org_k|)    covers the triangle test's own error, mapped through 1/|d_k| <= 50
// Rays with a direction component below 0.02 in magnitude are not filtered at all (their slab
// distances are too ill-conditioned for a fixed margin): leaf_margin() returns a negative value.
#ifndef ATN_LEAF_FILTER
#define ATN_LEAF_FILTER 1
#endif
__device__ __forceinline__ float leaf_margin(const V3& org, const V3& dir, const RayInv& ri)
{
    if (!ATN_LEAF_FILTER) return -1.0f;
    const float dmin = fminf(fminf(fabsf(dir.x), fabsf(dir.y)), fabsf(dir.z));
    if (!(dmin >= 0.02f)) return -1.0f;
    const float S = fabsf(ri.oxinvdir.x) + fabsf(ri.oxinvdir.y) + fabsf(ri.oxinvdir.z);
    const float M = fmaxf(fmaxf(fabsf(org.x), fabsf(org.y)), fabsf(org.z));
    return 1e-6f * S + 5e-4f * (1.0f + M);
}
__device__ __forceinline__ bool leaf_may_hit(float t0, float t1, float margin)
{
    // branch free; margin < 0 disables the filter
    const bool pass = t0 <= t1 + 2e-4f * (fabsf(t0) + fabsf(t1)) + margin;
    return pass | (margin < 0.0f);
}

// ---- ray sources / sinks of the traversal kernel -------------------------------------------
// A policy tells the kernels where ray i comes from and where its result goes:
//   load(i, ...)             ray i of the queue, its distance bound and early-exit mode (k_trace)
//   world(slot, ...)         the ray of a slot again (leaving an instance; k_exact)
//   t_max(slot)              its distance bound (k_exact)
//   finish(bv, slot, r, tie) end of the fast traversal: stores the result and returns true, or returns false when the
//                            result depends on a near tie the policy cares about (the slot then goes to the exact queue)
//   finish_hit(slot, h)      stores a resolved hit (used by finish)
//   finish_exact(slot, h)    the same for a hit the exact walk decided (k_exact, on the side stream)
__device__ __forceinline__ AtnHit resolved(const BvhView& bv, const TraceResult& r)
{
    AtnHit h;
    resolve_hit(bv, r, &h.objid, &h.primid, &h.mtrlid, &h.meshid);
    h.t = r.t; h.a = r.a; h.b = r.b; h.isVoxel = 0;
    return h;
}

struct ClosestPolicy {             // wavefront closest-hit (hitTest)
    static constexpr bool kNeedClass = false;
    PathBuffers pb;
    const int32_t* queue;          // path slots, nullptr = identity
    float t_max_all;
    int32_t* mesh_out;             // where to keep isect.meshid (first bounce only), or nullptr
    uint8_t pend_tag;              // bounce + 1 when the shade stage runs next to the exact walk (PathBuffers::pend), else 0
    __device__ __forceinline__ void load(int i, int* slot, V3* o, V3* d, float* t_max, int* mode, float* exit_dist) const
    {
        const int s = queue ? queue[i] : i;
        const float4 ro = pb.ray_o[s];
        const float4 rd = pb.ray_d[s];
        *slot = s; *o = V3(ro.x, ro.y, ro.z); *d = V3(rd.x, rd.y, rd.z);
        *t_max = t_max_all; *mode = EXIT_NONE; *exit_dist = 0.0f;
    }
    __device__ __forceinline__ void world(int slot, V3* o, V3* d) const
    {
        const float4 ro = pb.ray_o[slot];
        const float4 rd = pb.ray_d[slot];
        *o = V3(ro.x, ro.y, ro.z); *d = V3(rd.x, rd.y, rd.z);
    }
    __device__ __forceinline__ float t_max(int) const { return t_max_all; }
    __device__ __forceinline__ void finish_hit(int slot, const AtnHit& h) const
    {
        pb.hit0[slot] = make_float4(h.t, h.a, h.b, __int_as_float(h.primid));
        pb.hit1[slot] = make_int2(h.objid, h.mtrlid);
        if (mesh_out) mesh_out[slot] = h.meshid;
    }
    __device__ __forceinline__ void finish_exact(int slot, const AtnHit& h) const { finish_hit(slot, h); }
    __device__ __forceinline__ bool finish(const BvhView& bv, int slot, const TraceResult& r, const Tie& tie) const
    {
        if (tie.any) { if (pend_tag) pb.pend[slot] = pend_tag; return false; }     // the main surface stage leaves this path to the side pass
        finish_hit(slot, resolved(bv, r));
        return true;
    }
};

// Shadow rays: the result is a yes / no (HitShadowRay, pathtracing_impl.h:353-393), so a near tie only matters when the
// tied primitives could answer differently: another object or material (tie.cls), a material whose alpha can end the
// lookup, or a singular light whose distance falls inside the tie.
__device__ __forceinline__ bool shadow_tie_matters(const SceneView& sc, const AtnLightParameter& light, float dist, const AtnHit& h, const Tie& tie)
{
    if (h.objid < 0) return false;                               // nothing hit at all: the reference finds nothing either
    if (light.attrib & 2u) return false;                         // infinite light: any hit hides it
    if ((light.attrib & 1u) && h.t <= dist - tie.band) return false;   // singular light: something is clearly in front
    if (!tie.any) return false;
    if (tie.cls) return true;
    if ((light.attrib & 1u) && fabsf(h.t - dist) <= tie.band) return true;
    const AtnMaterialParameter& m = sc.materials[h.mtrlid < 0 ? 0 : h.mtrlid];
    return h.mtrlid >= 0 && (m.albedoMap >= 0 || m.baseColor.w < 1.0f);       // isTranslucentByAlpha looks at the tied hit's texel
}

// kRetrace: HitTestToTargetLight's loop (pathtracing_impl.h:288-348) for scenes with alpha blending on or a stencil material:
// a shadow ray whose closest hit is ignored (translucent by alpha / stencil) restarts behind it and joins the retrace queue,
// which the next pass of the same kernel traces (launch_bounces runs up to 9 such passes).  The per-ray loop state -- lookup
// number, stencil flag, the last object hit -- travels in the spare lane of the light-contribution record (sh_c.w).
__device__ __forceinline__ int32_t shadow_state_pack(int lookup, bool need_stencil, int32_t hitobj) { return lookup | (need_stencil ? 256 : 0) | ((hitobj + 1) << 9); }

template <bool kTap, bool kRetrace = false>
struct ShadowPolicy {              // wavefront shadow rays (hitShadowRay)
    static constexpr bool kNeedClass = !kRetrace;
    PathBuffers pb;
    SceneView sc;
    AtnBounceRecord* tap;
    int tap_cap, bounce;
    const int32_t* queue;          // path slots of this pass: q_shadow, or a retrace queue
    int32_t* q_retrace;            // kRetrace: where ignored hits go, and its counter
    int32_t* n_retrace;
    __device__ __forceinline__ void world(int slot, V3* o, V3* d) const
    {
        const float4 so = pb.sh_o[slot];
        const float4 sd = pb.sh_d[slot];
        // aten::ray(rayorg, raydir) renormalises the direction (pathtracing_impl.h:378); a retrace ray is built from that
        // direction with the same constructor, i.e. normalised once more (:330)
        *o = V3(so.x, so.y, so.z); *d = normalize(V3(sd.x, sd.y, sd.z));
        if (kRetrace && (__float_as_int(pb.sh_c[slot].w) & 255) != 0) *d = normalize(*d);
    }
    __device__ __forceinline__ float t_max(int slot) const { return pb.sh_o[slot].w - kEps; }
    // Area light: a shadow ray is visible iff its closest hit lies on the light's object (or nothing is hit).  Every point of
    // that object is inside its world bounds (SceneView::light_bounds), so an accepted hit clearly in front of the ray's
    // entry into those bounds settles it: the closest hit is that one or a nearer one, never the light.  Returns the
    // distance in front of which a hit ends the ray, -inf when the rule does not apply.
    __device__ __forceinline__ float area_exit_dist(int32_t light_id, const V3& o, const V3& d) const
    {
        if (!sc.light_bounds) return -kInf;
        const float4 lo = __ldg(reinterpret_cast<const float4*>(sc.light_bounds) + 2 * light_id);
        if (lo.w == 0.0f) return -kInf;
        const float4 hi = __ldg(reinterpret_cast<const float4*>(sc.light_bounds) + 2 * light_id + 1);
        if (!(fabsf(d.x) > 1e-12f && fabsf(d.y) > 1e-12f && fabsf(d.z) > 1e-12f)) return -kInf;
        const float ax = (lo.x - o.x) / d.x, bx = (hi.x - o.x) / d.x;
        const float ay = (lo.y - o.y) / d.y, by = (hi.y - o.y) / d.y;
        const float az = (lo.z - o.z) / d.z, bz = (hi.z - o.z) / d.z;
        const float t_in = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
        const float t_out = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
        if (!(t_in > 0.0f && t_in <= t_out)) return -kInf;
        // well clear of the bounds: 0.1 % of the distance and of the origin's magnitude (the bounds are padded already)
        return t_in - 1e-3f * (t_in + 1.0f + fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z)));
    }
    __device__ __forceinline__ void load(int i, int* slot, V3* o, V3* d, float* t_max, int* mode, float* exit_dist) const
    {
        const int s = queue[i];
        world(s, o, d);
        const float dist = pb.sh_o[s].w;
        const int32_t light_id = __float_as_int(pb.sh_d[s].w);
        const uint32_t attrib = sc.lights[light_id].attrib;
        *slot = s;
        *t_max = dist - kEps;                       // HitTestToTargetLight, pathtracing_impl.h:297-304
        // with the retrace loop the CLOSEST hit decides whether the ray goes on: no early exit
        *mode = kRetrace ? EXIT_NONE : ((attrib & 2u) ? EXIT_ANY_HIT : ((attrib & 1u) ? EXIT_HIT_BEFORE : EXIT_NONE));
        *exit_dist = dist;
        if (!kRetrace && (attrib & 3u) == 0u) {
            const float e = area_exit_dist(light_id, *o, *d);
            if (e > 0.0f) { *mode = EXIT_HIT_BEFORE; *exit_dist = e; }
        }
    }
    template <bool kLate>
    __device__ __forceinline__ void apply(int slot, const AtnHit& h) const
    {
        const float4 so = pb.sh_o[slot];
        const float4 sd = pb.sh_d[slot];
        const float4 c = pb.sh_c[slot];
        const AtnLightParameter& light = sc.lights[__float_as_int(sd.w)];
        Ray sr; sr.org = V3(so.x, so.y, so.z); sr.dir = normalize(V3(sd.x, sd.y, sd.z));
        bool vis;
        if (kRetrace) {
            const int32_t state = __float_as_int(c.w);
            const int lookup = state & 255;
            const bool need_stencil = (state & 256) != 0;
            int32_t hitobj = (state >> 9) - 1;
            const V3 orig_dir = sr.dir;
            if (lookup != 0) sr.dir = normalize(sr.dir);
            V3 next_org;
            const int res = shadow_lookup(sc, light, sr, orig_dir, so.w, h.objid, h.primid, h.a, h.b, h.t, h.mtrlid,
                                          lookup + 1 >= shadow_max_lookups(sc, need_stencil), need_stencil, &hitobj, &next_org);
            if (res == kShadowRetrace) {
                pb.sh_o[slot] = make_float4(next_org.x, next_org.y, next_org.z, so.w);
                pb.sh_c[slot] = make_float4(c.x, c.y, c.z, __int_as_float(shadow_state_pack(lookup + 1, need_stencil, hitobj)));
                q_retrace[atomicAdd(n_retrace, 1)] = slot;
                return;
            }
            vis = res == kShadowVisible;
        }
        else vis = shadow_visible(sc, light, sr, so.w, h.objid, h.primid, h.a, h.b, h.t, h.mtrlid);
        float4* dst = kLate ? pb.contrib_late : pb.contrib;
        if (vis) {
            float4 cc = dst[slot];
            cc.x += c.x; cc.y += c.y; cc.z += c.z;
            dst[slot] = cc;
        }
        if (kTap) {
            AtnBounceRecord& rec = tap[(size_t)slot * tap_cap + bounce];
            rec.shadow_visible = vis ? 1 : 0;
            if (vis) {
                const float4 cc = pb.contrib[slot]; const float4 l = pb.contrib_late[slot];
                rec.contrib[0] = cc.x + l.x; rec.contrib[1] = cc.y + l.y; rec.contrib[2] = cc.z + l.z;
            }
        }
    }
    __device__ __forceinline__ void finish_hit(int slot, const AtnHit& h) const { apply<false>(slot, h); }
    __device__ __forceinline__ void finish_exact(int slot, const AtnHit& h) const { apply<true>(slot, h); }
    __device__ __forceinline__ bool finish(const BvhView& bv, int slot, const TraceResult& r, const Tie& tie) const
    {
        const AtnHit h = resolved(bv, r);
        // (with the retrace loop every tie goes to the exact walk: which surface is the closest decides how the ray goes on)
        if (tie.any) {
            const int32_t light_id = __float_as_int(pb.sh_d[slot].w);
            const AtnLightParameter& light = sc.lights[light_id];
            bool matters = kRetrace || shadow_tie_matters(sc, light, pb.sh_o[slot].w, h, tie);
            if (matters && !kRetrace && (light.attrib & 3u) == 0u && h.objid >= 0) {
                // area light, and hit and runner-up both clearly in front of the light's bounds: hidden whichever is closest
                V3 o, d; world(slot, &o, &d);
                if (h.t <= area_exit_dist(light_id, o, d) - 2.0f * tie.band) matters = false;
            }
            if (matters) return false;
        }
        finish_hit(slot, h);
        return true;
    }
};

struct QueryClosestPolicy {        // atn_trace_closest
    static constexpr bool kNeedClass = false;
    const AtnRay* rays; const float* t_max_in; AtnHit* hits;
    __device__ __forceinline__ void world(int i, V3* o, V3* d) const { const AtnRay r = rays[i]; *o = V3(r.org); *d = V3(r.dir); }
    __device__ __forceinline__ float t_max(int i) const { return t_max_in ? fminf(t_max_in[i], kInf) : kInf; }
    __device__ __forceinline__ void load(int i, int* slot, V3* o, V3* d, float* tm, int* mode, float* exit_dist) const
    {
        world(i, o, d); *slot = i; *tm = t_max(i); *mode = EXIT_NONE; *exit_dist = 0.0f;
    }
    __device__ __forceinline__ void finish_hit(int i, const AtnHit& h) const { hits[i] = h; }
    __device__ __forceinline__ void finish_exact(int i, const AtnHit& h) const { hits[i] = h; }
    __device__ __forceinline__ bool finish(const BvhView& bv, int i, const TraceResult& r, const Tie& tie) const
    {
        if (tie.any) return false;
        hits[i] = resolved(bv, r);
        return true;
    }
};

struct QueryAnyPolicy {            // atn_trace_any: a singular light at distance dist[i]
    static constexpr bool kNeedClass = false;
    const AtnRay* rays; const float* dist; uint8_t* visible;
    __device__ __forceinline__ void world(int i, V3* o, V3* d) const { const AtnRay r = rays[i]; *o = V3(r.org); *d = V3(r.dir); }
    __device__ __forceinline__ float t_max(int i) const { return fminf(dist[i] - kEps, kInf); }
    __device__ __forceinline__ void load(int i, int* slot, V3* o, V3* d, float* tm, int* mode, float* exit_dist) const
    {
        world(i, o, d); *slot = i; *tm = t_max(i); *mode = EXIT_HIT_BEFORE; *exit_dist = dist[i];
    }
    __device__ __forceinline__ void finish_hit(int i, const AtnHit& h) const { visible[i] = (h.objid < 0 || h.t > dist[i]) ? 1 : 0; }
    __device__ __forceinline__ void finish_exact(int i, const AtnHit& h) const { finish_hit(i, h); }
    __device__ __forceinline__ bool finish(const BvhView&, int i, const TraceResult& r, const Tie& tie) const
    {
        // only a tie that straddles the light's distance can change the answer
        if (r.leaf >= 0 && tie.any && fabsf(r.t - dist[i]) <= tie.band) return false;
        visible[i] = (r.leaf < 0 || r.t > dist[i]) ? 1 : 0;
        return true;
    }
};

// ---------------------------------------------------------------------------------------------
// traversal kernel over QuadNodes (atn_bvh.h "Wide (4-slot) encoding")
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kStackStride = kTraceBlock * 8u;     // bytes between two levels of one lane's shared stack

// Two IEEE multiplications in one issue slot (FMUL2).  Only the products are packed: ptxas contracts a
// packed multiply feeding a packed add into FFMA2 even with .rn on both, which would round once where
// the reference rounds twice; the additions stay scalar add.rn.f32, which are never contracted.
__device__ __forceinline__ float2 mul2(float ax, float ay, float2 b)
{
    float2 r;
    asm("{\n\t.reg .b64 ra, rb, rc;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmul.rn.f32x2 rc, ra, rb;\n\tmov.b64 {%0, %1}, rc;\n\t}"
        : "=f"(r.x), "=f"(r.y) : "f"(ax), "f"(ay), "f"(b.x), "f"(b.y));
    return r;
}
// One 256-bit read-only load (LDG.E.256, sm_100): a lane's 32-byte sector in one L1 request instead of two.
#ifndef ATN_LDG256
#define ATN_LDG256 1
#endif
struct F4x2 { float4 lo, hi; };
__device__ __forceinline__ F4x2 ldg8(const float4* p)
{
    F4x2 r;
#if ATN_LDG256
    asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=f"(r.lo.x), "=f"(r.lo.y), "=f"(r.lo.z), "=f"(r.lo.w), "=f"(r.hi.x), "=f"(r.hi.y), "=f"(r.hi.z), "=f"(r.hi.w)
                 : "l"(p));
#else
    r.lo = ldg4(p); r.hi = ldg4(p + 1);
#endif
    return r;
}
// A whole QuadNode (7 x 16 bytes used) / a 48-byte leaf record, each as ONE asm statement: the loads are issued back to back.
__device__ __forceinline__ void ldg_node(const float4* p, F4x2& c0, F4x2& c1, F4x2& c2, F4x2& c3)
{
    asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%28];\n\t"
                 "ld.global.nc.v8.f32 {%8, %9, %10, %11, %12, %13, %14, %15}, [%28+32];\n\t"
                 "ld.global.nc.v8.f32 {%16, %17, %18, %19, %20, %21, %22, %23}, [%28+64];\n\t"
                 "ld.global.nc.v4.f32 {%24, %25, %26, %27}, [%28+96];"
                 : "=f"(c0.lo.x), "=f"(c0.lo.y), "=f"(c0.lo.z), "=f"(c0.lo.w), "=f"(c0.hi.x), "=f"(c0.hi.y), "=f"(c0.hi.z), "=f"(c0.hi.w),
                   "=f"(c1.lo.x), "=f"(c1.lo.y), "=f"(c1.lo.z), "=f"(c1.lo.w), "=f"(c1.hi.x), "=f"(c1.hi.y), "=f"(c1.hi.z), "=f"(c1.hi.w),
                   "=f"(c2.lo.x), "=f"(c2.lo.y), "=f"(c2.lo.z), "=f"(c2.lo.w), "=f"(c2.hi.x), "=f"(c2.hi.y), "=f"(c2.hi.z), "=f"(c2.hi.w),
                   "=f"(c3.lo.x), "=f"(c3.lo.y), "=f"(c3.lo.z), "=f"(c3.lo.w)
                 : "l"(p));
}
__device__ __forceinline__ void ldg_leaf(const float4* p, F4x2& c0, float4& c1)
{
    asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%12];\n\t"
                 "ld.global.nc.v4.f32 {%8, %9, %10, %11}, [%12+32];"
                 : "=f"(c0.lo.x), "=f"(c0.lo.y), "=f"(c0.lo.z), "=f"(c0.lo.w), "=f"(c0.hi.x), "=f"(c0.hi.y), "=f"(c0.hi.z), "=f"(c0.hi.w),
                   "=f"(c1.x), "=f"(c1.y), "=f"(c1.z), "=f"(c1.w)
                 : "l"(p));
}
__device__ __forceinline__ void sts64(uint32_t addr, int32_t x, int32_t y)
{
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" :: "r"(addr), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ int2 lds64(uint32_t addr)
{
    int2 v;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
    return v;
}

// Experiment switch: bit 0 = prefetch the subtree chosen by a node step, bit 1 = the one a pop returns.
#ifndef ATN_PREFETCH
#define ATN_PREFETCH 0
#endif
__device__ __forceinline__ void prefetch_ref(const BvhView& bv, int32_t ref, int32_t cur_inst)
{
    const float4* p = ref >= 0 ? bv.nodes + (size_t)ref * 8
                               : (cur_inst >= 0 ? bv.tris + (size_t)(~ref) * kTriStride : bv.insts + (size_t)(~ref) * kInstStride);
    asm volatile("prefetch.global.L1 [%0];" :: "l"(p));
}

// Entry distance and exit distance of the two slots held by one float4 triple of a QuadNode
// (aabb::hit, aabb.h:61-86: plane * invdir + oxinvdir, min / max per axis).
struct SlabPair { float t0a, t1a, t0b, t1b; };
__device__ __forceinline__ SlabPair slab_pair(const float4 mn_xy, const float4 mx_xy, const float4 z, const float2 inv_xy,
                                              const float2 inv_zz, const V3& ox)
{
    const float2 na = mul2(mn_xy.x, mn_xy.y, inv_xy), nb = mul2(mn_xy.z, mn_xy.w, inv_xy);
    const float2 fa = mul2(mx_xy.x, mx_xy.y, inv_xy), fb = mul2(mx_xy.z, mx_xy.w, inv_xy);
    const float2 nz = mul2(z.x, z.y, inv_zz), fz = mul2(z.z, z.w, inv_zz);
    const float nax = na.x + ox.x, nay = na.y + ox.y, naz = nz.x + ox.z;
    const float fax = fa.x + ox.x, fay = fa.y + ox.y, faz = fz.x + ox.z;
    const float nbx = nb.x + ox.x, nby = nb.y + ox.y, nbz = nz.y + ox.z;
    const float fbx = fb.x + ox.x, fby = fb.y + ox.y, fbz = fz.y + ox.z;
    SlabPair r;
    r.t1a = fminf(fminf(fmaxf(fax, nax), fmaxf(fay, nay)), fmaxf(faz, naz));
    r.t0a = fmaxf(fmaxf(fminf(fax, nax), fminf(fay, nay)), fminf(faz, naz));
    r.t1b = fminf(fminf(fmaxf(fbx, nbx), fmaxf(fby, nby)), fmaxf(fbz, nbz));
    r.t0b = fmaxf(fmaxf(fminf(fbx, nbx), fminf(fby, nby)), fminf(fbz, nbz));
    return r;
}

// compare-exchange of (key, reference) pairs, ascending keys
#define ATN_CE(ka, ra, kb, rb)                                          \
    {                                                                   \
        const bool sw_ = kb < ka;                                       \
        const float lo_ = fminf(ka, kb), hi_ = fmaxf(ka, kb);           \
        const int32_t rl_ = sw_ ? rb : ra, rh_ = sw_ ? ra : rb;         \
        ka = lo_; kb = hi_; ra = rl_; rb = rh_;                         \
    }

// Two-level closest-hit traversal with the reference's acceptance rules (threaded_bvh_traverser.h:98-304):
// inner boxes use aabb::hit, leaves are never box-tested, a triangle is accepted iff t_min < t && t < best
// (ties: lowest depth-first rank, which is what "first encountered" means in the reference's fixed
// visit order).
//
// Execution model: persistent warps.  Every lane owns one ray and a resumable traversal state; the
// warp advances all its rays one node / leaf per iteration and re-converges after each iteration.
// When fewer than kRefillThreshold lanes still have a ray, the warp goes back to the queue and the
// idle lanes pull new rays (one atomicAdd per warp, compacted with __ballot_sync/__popc), so
// incoherent rays do not leave most of the warp idle behind the slowest one.
//
// Node step: one 128-byte QuadNode (four 256-bit loads) decides up to four subtrees or leaves; the slots
// that are hit are ordered by entry distance with a 5-comparator network, the nearest is followed and
// the others go to the stack farthest first.  Every slot's box is tested with the reference's own slab
// test (a leaf's slot carries the box of the node whose hit makes the reference visit that leaf), so
// there is no approximate filter inside the trees; the only conservative pre-filter is the world-space
// box of an instance, tested before the ray is carried into the instance's space.
// Stack: kSmemStack entries per lane in shared memory (reference + box entry distance), spilling to
// local memory beyond that.  A popped subtree whose entry distance is already behind the current hit
// is dropped without fetching its node.
#ifndef ATN_TRACE_MIN_BLOCKS
#define ATN_TRACE_MIN_BLOCKS 8
#endif
// Which 32-byte chunks of a QuadNode are fetched through the TEXTURE pipe (tex1Dfetch of float4) instead of
// LDG.256; bit k = chunk k.  The kernel is co-limited by the LSU data pipe of L1 (one wavefront per distinct line
// and load instruction); the texture path reads the same cache through its own pipe, so moving part of the node
// there relieves the LSU one.  Measured on the bench workload (Mrays/s): 0: 1 749, 8: 1 809, 12: 1 807, 9: 1 814,
// 10: 1 821 (default), 14: 1 814, 15: 1 774; triangles through the texture pipe as well: no gain.
#ifndef ATN_NODE_TEX
#define ATN_NODE_TEX 10
#endif
// batching thresholds of the kStream form.  Measured on the 10 M-triangle scene (closest-hit Mrays/s, default 1 266):
// leaf / entry batch 1: 1 230, 3: 1 259; refill threshold 12: -6 %, 24: 1 285, 28: 1 256.  Also measured and not kept:
// part of the node through the texture pipe (748 - 1 165: the extra address registers make ptxas split the loads again),
// prefetch.global.L2 of the second-nearest subtree at push time (-12 ... -18 %: the LSU data pipe of L1, 80 % busy in
// this form, pays for every prefetch), 9 / 10 / 12 resident blocks with spills (-4 ... -50 %).
#ifndef ATN_STREAM_LEAF_BATCH
#define ATN_STREAM_LEAF_BATCH ATN_LEAF_BATCH
#endif
#ifndef ATN_STREAM_ENTRY_BATCH
#define ATN_STREAM_ENTRY_BATCH ATN_ENTRY_BATCH
#endif
#ifndef ATN_STREAM_REFILL
#define ATN_STREAM_REFILL ATN_REFILL_THRESHOLD
#endif
// Resident blocks per SM of the kStream form.  With 8 (64 registers) ptxas sinks half of a node's loads below the first
// use of the other half to save registers, and the node step waits for device memory twice; a larger register budget
// keeps the four loads together, which is worth more than the block it costs.  Closest-hit Mrays/s on the 10 M-triangle
// scene: 8 blocks 966 (784 when the first two loads are split as well), 7: 1 266, 6: 1 158, 5: 1 160.  The default form
// (L2-resident scenes) has the same split at 8 blocks but is issue bound: 7 or 6 blocks there are 1.5 % slower.
#ifndef ATN_STREAM_MIN_BLOCKS
#define ATN_STREAM_MIN_BLOCKS 7
#endif
// kCount: QuadNodes fetched and triangle tests run are counted (AtnStats.nodes_visited / tris_tested).  Only with
// atn_set_profiling(2): the two per-lane counters cost two registers of a 64-register kernel, 1.4 % of the bench frame.
template <class Policy, int kLocalStack, bool kStream, bool kCount = false>
__global__ void __launch_bounds__(kTraceBlock, kStream ? ATN_STREAM_MIN_BLOCKS : ATN_TRACE_MIN_BLOCKS)
k_trace(BvhView bv, Policy pol, const int32_t* __restrict__ count_ptr, int count_imm, int32_t* __restrict__ fetch, Visits* visits,
        int32_t* __restrict__ exact_q, int32_t* __restrict__ exact_count)
{
    __shared__ int2 s_stack[kSmemStack * kTraceBlock];   // (reference, entry distance bits) per lane and level
    __shared__ float s_world[12 * kTraceBlock];          // world-space org, dir, invdir, oxinvdir of the lane's ray
    // near-tie bookkeeping of the lane's ray (touched when a triangle is hit, an instance is entered, the ray ends):
    // band rho, band alpha, runner-up distance (any primitive), runner-up distance (other instance / material),
    // material of the best hit
    __shared__ float s_tie[6 * kTraceBlock];           // ..., [5] pathmax of the best hit's leaf
    int2 l_stack[kLocalStack];
    const int tid = threadIdx.x;
    const uint32_t stack_base = (uint32_t)__cvta_generic_to_shared(&s_stack[tid]);

    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int count = count_ptr ? *count_ptr : count_imm;
    const float t_min = bv.t_min;

    bool active = false, done = false, exhausted = false;
    int slot = 0, mode = EXIT_NONE, sp = 0, cur = kRefNull, cur_inst = -1;
    float exit_dist = 0.0f, t_cull = kInf;
    float cur_key = -kInf;            // largest box entry distance on the path from the root to `cur` (pathmax)
    V3 org, dir;
    RayInv ri;
    TraceResult best;
    best.t = kInf; best.a = 0.0f; best.b = 0.0f; best.leaf = -1; best.inst = -1;
    uint32_t n_nodes = 0, n_tris = 0;

    for (;;) {
        bool to_exact = false;
        if (done) {
            Tie tie{ false, false, 0.0f };
            if (exact_q && best.leaf >= 0) {
                const Band bd{ s_tie[0 * kTraceBlock + tid], s_tie[1 * kTraceBlock + tid] };
                tie.band = band_at(bd, best.t);
                const float t2 = s_tie[2 * kTraceBlock + tid];
                // a strictly nearest hit whose whole path starts no later than the runner-up is the reference's answer too
                const bool certain = best.t < t2 && s_tie[5 * kTraceBlock + tid] <= t2;
                tie.any = !certain && t2 <= best.t + tie.band;
                tie.cls = Policy::kNeedClass && tie.any && s_tie[3 * kTraceBlock + tid] <= best.t + tie.band;
            }
            to_exact = !pol.finish(bv, slot, best, tie);
            done = false;
        }
        if (exact_q) {           // rays the exact walk has to decide (rare): appended by the converged warp, one atomic per warp
            const unsigned em = __ballot_sync(full, to_exact);
            if (em != 0u) {
                const int leader = __ffs(em) - 1;
                int base = 0;
                if (lane == leader) base = atomicAdd(exact_count, __popc(em));
                base = __shfl_sync(full, base, leader);
                if (to_exact) exact_q[base + __popc(em & ((1u << lane) - 1u))] = slot;
            }
        }

        // ---------------- refill idle lanes ----------------
        const unsigned idle = __ballot_sync(full, !active);
        if (idle != 0u && !exhausted) {
            const int nidle = __popc(idle);
            const int leader = __ffs(idle) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(fetch, nidle);
            base = __shfl_sync(full, base, leader);
            if (base + nidle >= count) exhausted = true;
            if (!active) {
                const int i = base + __popc(idle & ((1u << lane) - 1u));
                if (i < count) {
                    float t_max;
                    pol.load(i, &slot, &org, &dir, &t_max, &mode, &exit_dist);
                    ri = ray_inv(org, dir);
                    s_world[0 * kTraceBlock + tid] = org.x; s_world[1 * kTraceBlock + tid] = org.y; s_world[2 * kTraceBlock + tid] = org.z;
                    s_world[3 * kTraceBlock + tid] = dir.x; s_world[4 * kTraceBlock + tid] = dir.y; s_world[5 * kTraceBlock + tid] = dir.z;
                    s_world[6 * kTraceBlock + tid] = ri.invdir.x; s_world[7 * kTraceBlock + tid] = ri.invdir.y; s_world[8 * kTraceBlock + tid] = ri.invdir.z;
                    s_world[9 * kTraceBlock + tid] = ri.oxinvdir.x; s_world[10 * kTraceBlock + tid] = ri.oxinvdir.y; s_world[11 * kTraceBlock + tid] = ri.oxinvdir.z;
                    t_cull = t_max;
                    best.t = kInf; best.a = 0.0f; best.b = 0.0f; best.leaf = -1; best.inst = -1;
#ifndef ATN_EXP_NO_BAND_LOAD
                    {
                        const Band bd = band_of(org, ri);
                        s_tie[0 * kTraceBlock + tid] = bd.rho; s_tie[1 * kTraceBlock + tid] = bd.alpha;
                        s_tie[2 * kTraceBlock + tid] = kInf; s_tie[3 * kTraceBlock + tid] = kInf;
                    }
#endif
                    sts64(stack_base, kRefDone, __float_as_int(-kInf));     // sentinel: popping it ends the ray
                    sp = 1; cur = bv.top_root; cur_inst = -1; cur_key = -kInf;
                    active = true;
                }
            }
        }
        if (__ballot_sync(full, active) == 0u) break;

        for (;;) {
            const unsigned act = __ballot_sync(full, active);
            if (act == 0u) break;
            if (!exhausted && __popc(act) < (kStream ? ATN_STREAM_REFILL : kRefillThreshold)) break;

            // ---- what every lane does in this iteration ----
            // kStream (scenes whose nodes and triangles do not fit the L2): the iteration is decided from the lanes'
            // current references alone and ALL its global loads -- a QuadNode, a TriLeaf or an InstLeafW per lane -- are
            // issued here, together, so that an iteration waits for device memory once.  The default form lets a lane
            // run node step, triangle test and instance entry back to back in one iteration (fewer iterations, which is
            // what counts when the set is L2-resident and the kernel is issue bound), at the price of up to three
            // dependent load latencies per iteration.
            const bool on_node = active && cur >= 0;
            F4x2 c0, c1, c2, c3;
            bool s_on_tri = false, s_on_entry = false;
            if (kStream) {
                const bool is_leaf = active && cur < 0 && cur > kRefDone;
                const bool w_tri = is_leaf && cur_inst >= 0, w_entry = is_leaf && cur_inst < 0;
                const unsigned mn = __ballot_sync(full, on_node);
                const unsigned mt = __ballot_sync(full, w_tri);
                const unsigned me = __ballot_sync(full, w_entry);
                const bool run_tris = (__popc(mt) >= ATN_STREAM_LEAF_BATCH) || (mn == 0u);
                s_on_tri = w_tri && run_tris;
                const bool busy = (mn != 0u) || (run_tris && mt != 0u);
                s_on_entry = w_entry && (__popc(me) >= ATN_STREAM_ENTRY_BATCH || !busy);
                const float4* p = on_node ? bv.nodes + (size_t)cur * 8
                                          : (s_on_tri ? bv.tris + (size_t)(~cur) * kTriStride : bv.insts + (size_t)(~cur) * kInstStride);
                // One asm statement per record kind, so that ptxas cannot sink part of a node's loads below the first use of
                // its other parts (it does, to save registers, and the node step then waits for device memory twice:
                // 966 vs 784 Mrays/s on the 10 M-triangle scene for two otherwise identical builds).
                if (on_node) ldg_node(p, c0, c1, c2, c3);
                else if (s_on_tri || s_on_entry) ldg_leaf(p, c0, c1.lo);
            }

            // ---- phase A: QuadNode ----
            if (on_node) {
                if (!kStream) {
                    const float4* np = bv.nodes + (size_t)cur * 8;
                    // chunk k through the texture pipe (two 128-bit fetches; one for the last chunk, whose upper half is
                    // padding) or through LDG.256, see ATN_NODE_TEX
                    if (ATN_NODE_TEX & 1) { c0.lo = tex1Dfetch<float4>(bv.node_tex, cur * 8 + 0); c0.hi = tex1Dfetch<float4>(bv.node_tex, cur * 8 + 1); }
                    else c0 = ldg8(np + 0);
                    if (ATN_NODE_TEX & 2) { c1.lo = tex1Dfetch<float4>(bv.node_tex, cur * 8 + 2); c1.hi = tex1Dfetch<float4>(bv.node_tex, cur * 8 + 3); }
                    else c1 = ldg8(np + 2);
                    if (ATN_NODE_TEX & 4) { c2.lo = tex1Dfetch<float4>(bv.node_tex, cur * 8 + 4); c2.hi = tex1Dfetch<float4>(bv.node_tex, cur * 8 + 5); }
                    else c2 = ldg8(np + 4);
                    if (ATN_NODE_TEX & 8) c3.lo = tex1Dfetch<float4>(bv.node_tex, cur * 8 + 6);
                    else c3 = ldg8(np + 6);
                }
                const float4 a0 = c0.lo, a1 = c0.hi, a2 = c1.lo, b0 = c1.hi, b1 = c2.lo, b2 = c2.hi;
                const int4 rf = make_int4(__float_as_int(c3.lo.x), __float_as_int(c3.lo.y), __float_as_int(c3.lo.z), __float_as_int(c3.lo.w));
                const float2 inv_xy = make_float2(ri.invdir.x, ri.invdir.y);
                const float2 inv_zz = make_float2(ri.invdir.z, ri.invdir.z);
                const SlabPair pa = slab_pair(a0, a1, a2, inv_xy, inv_zz, ri.oxinvdir);
                const SlabPair pb = slab_pair(b0, b1, b2, inv_xy, inv_zz, ri.oxinvdir);
                if (kCount) n_nodes++;
                // t0 = max(entry, t_min) <= min(exit, running cull bound); empty slots carry an infinite box and never
                // pass (t_cull is finite: every source clamps it to kInf = FLT_MAX)
                float k0 = fmaxf(pa.t0a, t_min), k1 = fmaxf(pa.t0b, t_min), k2 = fmaxf(pb.t0a, t_min), k3 = fmaxf(pb.t0b, t_min);
                int32_t r0 = rf.x, r1 = rf.y, r2 = rf.z, r3 = rf.w;
                const bool h0 = k0 <= fminf(pa.t1a, t_cull);
                const bool h1 = k1 <= fminf(pa.t1b, t_cull);
                const bool h2 = k2 <= fminf(pb.t1a, t_cull);
                const bool h3 = k3 <= fminf(pb.t1b, t_cull);
                const int nhit = (int)h0 + (int)h1 + (int)h2 + (int)h3;
                // (taking the path maximum only for the slots that are kept -- the followed one and the pushed ones -- instead
                // of all four measured the same: 1 806 vs 1 807 Mrays/s)
                k0 = h0 ? fmaxf(k0, cur_key) : kInf; k1 = h1 ? fmaxf(k1, cur_key) : kInf;       // pathmax keys
                k2 = h2 ? fmaxf(k2, cur_key) : kInf; k3 = h3 ? fmaxf(k3, cur_key) : kInf;
                ATN_CE(k0, r0, k1, r1) ATN_CE(k2, r2, k3, r3) ATN_CE(k0, r0, k2, r2) ATN_CE(k1, r1, k3, r3) ATN_CE(k1, r1, k2, r2)
                cur = nhit ? r0 : kRefNull;
                cur_key = k0;
#if ATN_PREFETCH & 1
                // the chosen subtree's record is fetched at the top of the next iteration, after the triangle / entry / pop
                // phases of the other lanes: start moving it towards the L1 now
                if (!kStream && nhit) prefetch_ref(bv, cur, cur_inst);
#endif
                const int np_ = nhit - 1;                 // entries left behind, nearest on top
                if (np_ > 0) {
                    if (sp + np_ <= kSmemStack) {
                        const uint32_t top = stack_base + (uint32_t)(sp + np_) * kStackStride;
                        sts64(top - kStackStride, r1, __float_as_int(k1));
                        if (np_ > 1) sts64(top - 2u * kStackStride, r2, __float_as_int(k2));
                        if (np_ > 2) sts64(top - 3u * kStackStride, r3, __float_as_int(k3));
                    }
                    else {
                        const int32_t rr[3] = { r1, r2, r3 };
                        const float kk[3] = { k1, k2, k3 };
#pragma unroll
                        for (int k = 2; k >= 0; k--) {
                            if (k < np_) {
                                const int at = sp + (np_ - 1 - k);
                                if (at < kSmemStack) sts64(stack_base + (uint32_t)at * kStackStride, rr[k], __float_as_int(kk[k]));
                                else l_stack[at - kSmemStack] = make_int2(rr[k], __float_as_int(kk[k]));
                            }
                        }
                    }
                    sp += np_;
                }
            }

            // ---- phase B: triangle leaf (triangle::hit, triangle.h:40-67) ----
            bool on_tri = s_on_tri;
            if (!kStream) {
                const bool want_tri = active && cur < 0 && cur > kRefDone && cur_inst >= 0;       // a leaf, not one of the reserved values
                bool run_tris = true;
                if (kLeafBatch > 1) {
                    const unsigned mt = __ballot_sync(full, want_tri);
                    const unsigned mn = __ballot_sync(full, on_node);
                    run_tris = (__popc(mt) >= kLeafBatch) || (mn == 0u);
                }
                on_tri = want_tri && run_tris;
            }
            bool finished = false;
            if (on_tri) {
                const int32_t leaf = ~cur;
                if (!kStream) {
                    const float4* tp = bv.tris + (size_t)leaf * kTriStride;
                    c0 = ldg8(tp);
                    c1.lo = ldg4(tp + 2);
                }
                const float4 a = c0.lo, b = c0.hi;
                const float4 c = c1.lo;
                const TriHit h = intersect_triangle_edges(org, dir, V3(a.x, a.y, a.z), V3(b.x, b.y, b.z), V3(c.x, c.y, c.z));
                if (kCount) n_tris++;
                if (h.hit && h.t < kInf && t_min < h.t) {
#if ATN_HIT_BRANCHES
                    // Three cases, as branches: only a couple of lanes of a warp are here at a time, so skipped instructions
                    // are saved issue slots.  (1) The ray's first hit: nothing to compare with.  (2) A hit beyond the cull
                    // bound of the best one (its box survived, the triangle is farther): it neither wins nor ties.  (3) The
                    // general case below.
                    if (best.leaf < 0) {
                        const Band bd{ s_tie[0 * kTraceBlock + tid], s_tie[1 * kTraceBlock + tid] };
                        if (mode == EXIT_ANY_HIT || (mode == EXIT_HIT_BEFORE && h.t <= exit_dist - band_at(bd, h.t))) finished = true;
                        best.t = h.t; best.a = h.a; best.b = h.b; best.leaf = leaf; best.inst = cur_inst;
                        t_cull = cull_bound(bd, h.t);
                        s_tie[5 * kTraceBlock + tid] = cur_key;
                        if (Policy::kNeedClass) s_tie[4 * kTraceBlock + tid] = b.w;        // material of the best hit
                    }
                    else if (h.t > t_cull) {}
                    else
#endif
                    {
                    const bool closer = h.t < best.t ||
                        (h.t == best.t && (cur_inst < best.inst || (cur_inst == best.inst && leaf < best.leaf)));
                    // Runner-up bookkeeping (see "near ties"): only a hit within the cull bound of the best one can tie with
                    // it -- t_cull is already in a register, so the common cases (a far hit that loses, a first hit) cost one
                    // comparison; the shared-memory update below runs for actual near ties only.
                    float t2 = kInf;                           // distance of the hit that is NOT the best afterwards, if it can tie
                    float nc = t_cull;
                    // another SBVH leaf of the best primitive gives the same t, a, b bit for bit: not a runner-up
                    const bool same = best.leaf >= 0 && h.t == best.t && h.a == best.a && h.b == best.b;
                    if (closer) {
                        const Band bd{ s_tie[0 * kTraceBlock + tid], s_tie[1 * kTraceBlock + tid] };
                        nc = cull_bound(bd, h.t);
                        if (best.leaf >= 0 && best.t <= nc) t2 = best.t;
                        if (mode == EXIT_ANY_HIT || (mode == EXIT_HIT_BEFORE && h.t <= exit_dist - band_at(bd, h.t))) finished = true;
                    }
                    else if (h.t <= t_cull) t2 = h.t;
#ifndef ATN_EXP_NO_T2
                    if (same) s_tie[5 * kTraceBlock + tid] = fminf(s_tie[5 * kTraceBlock + tid], cur_key);   // the easier path to the best primitive counts
                    if (t2 < kInf) {
                        if (!same) {
                            s_tie[2 * kTraceBlock + tid] = fminf(s_tie[2 * kTraceBlock + tid], t2);
                            if (Policy::kNeedClass && (cur_inst != best.inst || __float_as_int(b.w) != __float_as_int(s_tie[4 * kTraceBlock + tid])))
                                s_tie[3 * kTraceBlock + tid] = fminf(s_tie[3 * kTraceBlock + tid], t2);
                        }
                    }
#endif
                    if (closer) {
                        best.t = h.t; best.a = h.a; best.b = h.b; best.leaf = leaf; best.inst = cur_inst;
                        t_cull = nc;
                        if (!same) s_tie[5 * kTraceBlock + tid] = cur_key;
                        if (Policy::kNeedClass) s_tie[4 * kTraceBlock + tid] = b.w;        // material of the best hit
                    }
                    }
                }
                cur = kRefNull;
            }

            // ---- phase E: instance entry (threaded_bvh_traverser.h:146-185), batched ----
            bool want_entry = s_on_entry;
            unsigned me = 0u;
            if (!kStream) {
                want_entry = active && cur < 0 && cur > kRefDone && cur_inst < 0;
                me = __ballot_sync(full, want_entry);
            }
            else me = __ballot_sync(full, s_on_entry);
            if (me != 0u) {
                const unsigned busy = kStream ? 0u : __ballot_sync(full, on_node || on_tri);
                if (kStream || __popc(me) >= kEntryBatch || busy == 0u) {
                    if (want_entry) {
                        const int32_t leaf = ~cur;
                        if (!kStream) {
                            const float4* ip = bv.insts + (size_t)leaf * kInstStride;
                            c0 = ldg8(ip);
                            c1.lo = ldg4(ip + 2);
                        }
                        const F4x2 iw = c0;
                        const int4 il = make_int4(__float_as_int(iw.lo.x), __float_as_int(iw.lo.y), __float_as_int(iw.lo.z), __float_as_int(iw.lo.w));
                        const float4 wmin = iw.hi, wmax = c1.lo;
                        cur = kRefNull;
                        // conservative pre-filter (see leaf_margin): does the world ray come near the instance at all?
                        float t0w, t1w;
                        slab_test(ri, wmin.x, wmax.x, wmin.y, wmax.y, wmin.z, wmax.z, t_min, kInf, &t0w, &t1w);
                        const float far = (best.leaf >= 0) ? t_cull : kInf;
                        const bool may = leaf_may_hit(t0w, fminf(t1w, far), leaf_margin(org, dir, ri));
                        if (il.y != kRefNull && may) {
                            if (sp < kSmemStack) sts64(stack_base + (uint32_t)sp * kStackStride, kRefExit, __float_as_int(-kInf));
                            else l_stack[sp - kSmemStack] = make_int2(kRefExit, __float_as_int(-kInf));
                            sp++;
                            cur_inst = leaf;
                            if (il.z >= 0) {
                                const AtnMat4 m = bv.matrices[il.z];
                                const Ray tr = make_ray(mat_apply(m, org), mat_apply_xyz(m, dir));   // mat4::applyRay
                                org = tr.org; dir = tr.dir;
                                ri = ray_inv(org, dir);
                                // the band covers every space the ray has been carried into; the cull bound follows it
                                const Band nb = band_of(org, ri);
                                const Band bd{ fmaxf(s_tie[0 * kTraceBlock + tid], nb.rho), fmaxf(s_tie[1 * kTraceBlock + tid], nb.alpha) };
                                s_tie[0 * kTraceBlock + tid] = bd.rho; s_tie[1 * kTraceBlock + tid] = bd.alpha;
                                if (best.leaf >= 0) t_cull = cull_bound(bd, best.t);
                            }
                            cur = il.y;
                        }
                    }
                }
            }

            // ---- phase C: pop.  Runs for a few lanes at a time, so it is kept short: entries whose subtree starts behind
            // the current hit are skipped in a three-instruction loop, the sentinel at the bottom of the stack (entry
            // distance -inf, like the exit markers) ends the ray without a separate emptiness test. ----
            if (active && !finished && cur == kRefNull) {
#pragma unroll 1
                for (;;) {
                    int2 e;
#pragma unroll 1
                    do {
                        sp--;
                        e = (sp < kSmemStack) ? lds64(stack_base + (uint32_t)sp * kStackStride) : l_stack[sp - kSmemStack];
                    } while (__int_as_float(e.y) > t_cull);
                    cur = e.x; cur_key = __int_as_float(e.y);
#if ATN_PREFETCH & 2
                    if (!kStream && cur > kRefDone) prefetch_ref(bv, cur, cur_inst);
#endif
                    if (cur != kRefExit) break;
                    // back to the top layer with the world ray (threaded_bvh_traverser.h:290-300), then pop again
                    cur_inst = -1;
                    org = V3(s_world[0 * kTraceBlock + tid], s_world[1 * kTraceBlock + tid], s_world[2 * kTraceBlock + tid]);
                    dir = V3(s_world[3 * kTraceBlock + tid], s_world[4 * kTraceBlock + tid], s_world[5 * kTraceBlock + tid]);
                    ri.invdir = V3(s_world[6 * kTraceBlock + tid], s_world[7 * kTraceBlock + tid], s_world[8 * kTraceBlock + tid]);
                    ri.oxinvdir = V3(s_world[9 * kTraceBlock + tid], s_world[10 * kTraceBlock + tid], s_world[11 * kTraceBlock + tid]);
                }
                finished = (cur == kRefDone);
            }
            if (finished) { active = false; done = true; }
        }
    }
    if (kCount) {
        for (int o = 16; o > 0; o >>= 1) {
            n_nodes += __shfl_xor_sync(full, n_nodes, o);
            n_tris += __shfl_xor_sync(full, n_tris, o);
        }
        if (lane == 0 && visits) {
            atomicAdd(&visits->nodes, (unsigned long long)n_nodes);
            atomicAdd(&visits->tris, (unsigned long long)n_tris);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// The exact walk: ThreadedBvhTraverser<true>::Traverse<Closest> (threaded_bvh_traverser.h:98-304) as it stands -- the
// reference's own node lists, its stackless hit / miss links, its depth-first order, its running t_max through the biased
// slab test -- for the few rays whose answer depends on that order (see "near ties" above).  One thread per ray; the
// queue holds a fraction of a percent of a wavefront, so this kernel is a short tail behind k_trace, not a second
// traversal of the frame.
// ---------------------------------------------------------------------------------------------
// The walk is written as a state machine with ONE round of loads per iteration whatever the lane is doing -- a node
// (3 x 16 B), an object record, a matrix (3 rows), a triangle record, three vertex positions -- so that the lanes of a warp,
// which are at unrelated nodes of unrelated rays, wait for memory once per iteration instead of once per kind of step.
__device__ __forceinline__ AtnHit ref_traverse(const RefView& rv, const V3& worg, const V3& wdir, float t_max)
{
    enum : int { PH_NODE, PH_OBJECT, PH_MATRIX, PH_TRIANGLE, PH_POSITIONS };
    AtnHit isect;
    isect.t = kInf; isect.objid = -1; isect.mtrlid = -1; isect.meshid = -1; isect.primid = -1; isect.a = 0.0f; isect.b = 0.0f; isect.isVoxel = 0;
    const float4* const top = reinterpret_cast<const float4*>(rv.nodes + rv.offsets[0]);
    const float4* list = top;
    V3 org = worg, dir = wdir;
    RayInv ri = ray_inv(org, dir);
    int32_t objid = -1, meshid = -1, top_hit = -1, top_miss = -1;
    int32_t link_hit = 0, link_miss = 0, exid = 0, primid = 0, mtrlid = 0, tri_mesh = 0;
    int phase = PH_NODE;
    const float4* a0 = list; const float4* a1 = list + 1; const float4* a2 = list + 2;      // node 0 of the top layer
    for (;;) {
        const float4 v0 = __ldg(a0), v1 = __ldg(a1), v2 = __ldg(a2);
        // set when this iteration finishes a node: the reference's `is_hit`, or an entry into a bottom tree (root = node 0)
        bool decided = false, is_hit = false, entered = false;
        if (phase == PH_NODE) {
            link_hit = (int32_t)v0.w; link_miss = (int32_t)v1.w;
            if (v2.x >= 0 || v2.y >= 0) {                               // ThreadedBvhNode::isLeaf
                if (v2.z >= 0) {                                        // a top-layer leaf with a bottom tree
                    objid = (int32_t)v2.x; meshid = (int32_t)v2.w;
                    exid = (int32_t)((uint32_t)__float_as_int(v2.z) & 0x7fffu);
                    top_hit = link_hit; top_miss = link_miss;
                    phase = PH_OBJECT;
                    a0 = a1 = a2 = reinterpret_cast<const float4*>(rv.objects + objid);        // type, area, object_id, mtx_id
                }
                else if (v2.y >= 0) {                                   // a triangle
                    primid = (int32_t)v2.y;
                    phase = PH_TRIANGLE;
                    a0 = reinterpret_cast<const float4*>(rv.triangles + primid); a1 = a0 + 1; a2 = a1;
                }
                else decided = true;                                    // a leaf that holds nothing traversable (sphere): is_hit stays false
            }
            else {
                float t0;
                is_hit = aabb_hit(ri, V3(v0.x, v0.y, v0.z), V3(v1.x, v1.y, v1.z), rv.t_min, t_max, &t0);
                decided = true;
            }
        }
        else if (phase == PH_OBJECT) {
            const int32_t mtx_id = __float_as_int(v0.w);
            if (mtx_id >= 0) {
                phase = PH_MATRIX;
                a0 = reinterpret_cast<const float4*>(rv.matrices + mtx_id + 1); a1 = a0 + 1; a2 = a0 + 2;    // rows 0-2 of W2L
            }
            else {
                org = worg; dir = wdir; ri = ray_inv(org, dir);
                list = reinterpret_cast<const float4*>(rv.nodes + rv.offsets[exid]);
                entered = true;
            }
        }
        else if (phase == PH_MATRIX) {
            AtnMat4 m;
            m.m[0] = v0.x; m.m[1] = v0.y; m.m[2] = v0.z; m.m[3] = v0.w; m.m[4] = v1.x; m.m[5] = v1.y; m.m[6] = v1.z; m.m[7] = v1.w;
            m.m[8] = v2.x; m.m[9] = v2.y; m.m[10] = v2.z; m.m[11] = v2.w; m.m[12] = m.m[13] = m.m[14] = 0.0f; m.m[15] = 1.0f;
            const Ray tr = make_ray(mat_apply(m, worg), mat_apply_xyz(m, wdir));       // mat4::applyRay
            org = tr.org; dir = tr.dir; ri = ray_inv(org, dir);
            list = reinterpret_cast<const float4*>(rv.nodes + rv.offsets[exid]);
            entered = true;
        }
        else if (phase == PH_TRIANGLE) {
            mtrlid = __float_as_int(v1.z); tri_mesh = __float_as_int(v1.w);
            phase = PH_POSITIONS;
            a0 = reinterpret_cast<const float4*>(rv.positions + __float_as_int(v0.x));
            a1 = reinterpret_cast<const float4*>(rv.positions + __float_as_int(v0.y));
            a2 = reinterpret_cast<const float4*>(rv.positions + __float_as_int(v0.z));
        }
        else {                                                          // PH_POSITIONS: triangle::hit (triangle.h:40-67)
            const TriHit h = intersect_triangle(org, dir, V3(v0.x, v0.y, v0.z), V3(v1.x, v1.y, v1.z), V3(v2.x, v2.y, v2.z));
            is_hit = h.hit && h.t < kInf;
            decided = true;
            const float tt = is_hit ? h.t : kInf;
            if (rv.t_min < tt && is_hit && tt < isect.t) {
                isect.t = h.t; isect.a = h.a; isect.b = h.b;
                isect.objid = objid; isect.primid = primid; isect.mtrlid = mtrlid;
                isect.meshid = tri_mesh < 0 ? meshid : tri_mesh;
                t_max = isect.t;
            }
        }
        if (decided || entered) {
            int32_t next = entered ? 0 : (is_hit ? link_hit : link_miss);
            if (next < 0) {                                             // end of a list (threaded_bvh_traverser.h:290-300)
                next = is_hit ? top_hit : top_miss;                     // both -1 while in the top layer: the walk ends
                top_hit = -1; top_miss = -1;
                list = top; org = worg; dir = wdir; ri = ray_inv(org, dir);
                if (next < 0) break;
            }
            phase = PH_NODE;
            a0 = list + (size_t)next * 3; a1 = a0 + 1; a2 = a0 + 2;
        }
    }
    return isect;
}

template <class Policy>
__global__ void __launch_bounds__(128)
k_exact(RefView rv, Policy pol, const int32_t* __restrict__ exact_q, const int32_t* __restrict__ exact_count)
{
#ifdef ATN_EXP_EXACT_NOOP
    const int count = 0;      // timing experiment: what the exact walk costs beyond its own work
#else
    const int count = *exact_count;
#endif
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const int slot = exact_q[i];
        V3 o, d;
        pol.world(slot, &o, &d);
        pol.finish_exact(slot, ref_traverse(rv, o, d, pol.t_max(slot)));
    }
}

