// atn_kernels.cu -- sm_100a kernels of the wavefront NEE path tracer + the C ABI (include/aten_b200.h).
//
// Pipeline per camera sample (replaces reference src/libidaten/kernel/pathtracing.cpp:91-153):
//
//   k_generate              genPath            (pt_standard_impl.cu:25-50, pathtracing_impl.h:65-110)
//   for bounce:
//     k_trace<ClosestPolicy> hitTest           (pt_standard_impl.cu:57-229, threaded_bvh_traverser.h:98-304)
//     k_surface             shade, first part + ShadeMiss (pathtracing_impl.cu:18-238, pt_standard_impl.cu:330-362;
//                                               HOST ordering, pathtracing.cpp:91-236)
//     k_nee                 FillShadowRay / SampleLight / ComputeRadianceNEE (pathtracing_impl.h:178-264)
//     k_bsdf                Russian roulette, sampleMaterial, PrepareForNextBounce (pathtracing_impl.h:680-743)
//     k_trace<ShadowPolicy>  hitShadowRay       (pathtracing_impl.cu:240-280, pathtracing_impl.h:266-393)
//   k_gather                gather             (pathtracing_impl.cu:282-324, film.cpp:61-71)
//
// B200-native choices (DESIGN.md has the numbers):
//   * path state is SoA float4 / u32 arrays indexed by path slot; there is no AoS PathThroughput;
//   * the reference's scan-based stream compaction (StreamCompaction.cu, >= 7 launches + syncs per
//     bounce) is replaced by device-side queues appended with warp-aggregated atomics
//     (__ballot_sync / __popc / one atomicAdd per warp); the host never reads a count, so a frame
//     is a fixed launch sequence with no synchronisation;
//   * traversal kernels are persistent and warp-synchronous: grid = SMs x resident blocks, every lane
//     owns a resumable ray, the warp refills idle lanes from the queue with one atomicAdd;
//   * nodes are 128-byte 4-slot records (an exact collapse of the reference's binary trees, atn_bvh.h)
//     fetched with 256-bit read-only loads (LDG.E.256), triangles are 64-byte gathered records; the
//     traversal stack is a per-lane shared-memory stack of (reference, entry distance) pairs that
//     spills to local memory beyond kSmemStack;
//   * no tensor cores: the path is pointer chasing + FP32 slab / Moeller-Trumbore arithmetic.
// Compiled with -fmad=false: the intersection arithmetic must round exactly like the reference's
// x86 build for primitive ids to match bit for bit.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdarg.h>
#include <stdlib.h>
#include <mutex>
#include <string>
#include <vector>
#include <algorithm>

#include "atn_math.cuh"
#include "atn_svgf.cuh"
#include "atn_restir.cuh"
#include "atn_bvh.h"
#include "atn_cache.h"

using namespace atn;

// ---------------------------------------------------------------------------------------------
// error handling
// ---------------------------------------------------------------------------------------------
static thread_local std::string g_last_error;

static int set_error(int code, const char* fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

#define CUDA_TRY(expr)                                                                              \
    do {                                                                                            \
        cudaError_t e__ = (expr);                                                                   \
        if (e__ != cudaSuccess)                                                                     \
            return set_error(ATN_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), \
                             __FILE__, __LINE__);                                                   \
    } while (0)

// ---------------------------------------------------------------------------------------------
// device-side views
// ---------------------------------------------------------------------------------------------
constexpr int kInstStride = 4;                   // 16-byte words per InstLeafW
constexpr int kTriStride = 4;                    // 16-byte words per TriLeaf

struct BvhView {
    const float4* nodes;     // QuadNode as 8 x float4
    const float4* tris;      // TriLeaf as 4 x float4 (the last one is padding)
    const float4* insts;     // InstLeafW as 4 x 16 B
    const AtnMat4* matrices;
    int32_t top_root;
    float t_min;
    cudaTextureObject_t node_tex;   // the QuadNode array again, as a linear float4 texture (see ATN_NODE_TEX)
};

// The reference's own threaded node lists (48-byte nodes, hit / miss links), kept on the device for the exact walk that
// decides near ties (k_exact): all lists in one array, list k starting at offsets[k].
struct RefView {
    const AtnBvhNode* nodes;
    const uint32_t* offsets;
    const AtnObjectParameter* objects;
    const AtnMat4* matrices;
    const AtnTriangleParameter* triangles;
    const AtnVec4* positions;
    float t_min;
};

constexpr int kMaxBounce = 16;
#ifndef ATN_SMEM_STACK
#define ATN_SMEM_STACK 14
#endif
constexpr int kSmemStack = ATN_SMEM_STACK;     // shared-memory stack entries per thread
// overflow of the shared-memory stack, local memory: two instantiations of the kernel, picked per scene from the
// encoder's exact worst-case depth (a larger local frame costs ~1 % even when it is never touched)
constexpr int kLocalStackSmall = 48, kLocalStackLarge = 240;
constexpr int kTraceBlock = 128;
#ifndef ATN_HIT_BRANCHES
#define ATN_HIT_BRANCHES 1
#endif
#ifndef ATN_REFILL_THRESHOLD
#define ATN_REFILL_THRESHOLD 20
#endif
constexpr int kRefillThreshold = ATN_REFILL_THRESHOLD;   // refill a warp when fewer lanes than this still own a ray
#ifndef ATN_LEAF_BATCH
#define ATN_LEAF_BATCH 6
#endif
constexpr int kLeafBatch = ATN_LEAF_BATCH;               // lanes that must hold a triangle leaf before the test runs
#ifndef ATN_ENTRY_BATCH
#define ATN_ENTRY_BATCH 6
#endif
constexpr int kEntryBatch = ATN_ENTRY_BATCH;             // lanes that must want to enter an instance before the transform runs

// Per-frame device counters; zeroed with one memset per sample pass.
struct Counters {
    int32_t n_active[kMaxBounce + 1];   // rays queued for closest-hit at bounce b
    int32_t n_lit[kMaxBounce];          // surface hits that go on to NEE + BSDF sampling at bounce b
    int32_t n_shadow[kMaxBounce];       // shadow rays queued at bounce b
    int32_t fetch_closest[kMaxBounce];  // persistent-thread work cursors
    int32_t fetch_shadow[kMaxBounce];
    int32_t fetch_query;
    int32_t n_exact_closest[kMaxBounce];  // rays handed to the exact walk (near ties), per bounce
    int32_t n_exact_shadow[kMaxBounce];
    int32_t n_exact_query;
    int32_t fetch_shade[kMaxBounce][4]; // work cursors of the shade kernels of bounce b (surface, nee, bsdf)
    // rays that restart behind an ignored hit (alpha / stencil), per bounce and lookup: queue length, work cursor of the pass
    // that traces them, and their share of the exact walk
    struct Retrace { int32_t n[10], fetch[10], n_exact[10]; } retrace[kMaxBounce];
    int32_t pad[2];
};

// Ray totals of one atn_render call (not cleared per sample).
struct Totals { unsigned long long closest, shadow, exact; };
// Traversal work of one call: QuadNodes fetched and triangle tests run (all lanes, all traversal kernels).
struct Visits { unsigned long long nodes, tris; };

struct PathBuffers {
    float4* ray_o;        // org.xyz, -
    float4* ray_d;        // dir.xyz, -
    float4* hit0;         // t, a, b, bits(primid)
    int2* hit1;           // objid, mtrlid   (objid < 0: miss)
    float4* throughput;   // xyz, pdfb
    float4* contrib;      // xyz, samples
    float4* contrib_late; // light contributions of the shadow rays the exact walk decided (k_exact); zero for all other paths.
                          // They arrive on the side stream at any time before the gather, which adds them: the main stream
                          // never waits for them, and a path's sum does not depend on when they arrived
    uint32_t* sampler_dim;
    uint32_t* flags;      // bit0 terminated, bit1 singular
    int32_t* queue[2];    // path slots to trace at bounce b (b & 1)
    int32_t* q_lit;       // path slots that passed the surface stage
    int32_t* q_shadow;    // path slots with a shadow ray to trace
    int32_t* q_exact;     // path slots whose closest-hit traversal ended in a near tie: the exact walk decides them (k_exact)
    int32_t* q_exact_s;   // the same for shadow rays (its own list: the two exact walks of a bounce overlap other work)
    int32_t* q_retrace[2]; // shadow rays restarting behind an ignored hit (alpha / stencil): ping-pong between lookups
    uint8_t* pend;        // bounce + 1 of the bounce whose closest hit the exact walk decides (0: none since k_generate).  Written
                          // by the traversal, read by the main surface stage, never touched by the side pass -- which
                          // overwrites hit0 / hit1 of such a path while the main surface stage is still running
    float4* surf0;        // Surface: p.xyz, u
    float4* surf1;        //          orienting normal.xyz, v
    float4* surf2;        //          albedo.xyz, bits(mtrlid)
    float4* sh_o;         // shadow org.xyz, dist            (indexed by path slot)
    float4* sh_d;         // shadow dir.xyz, bits(light id)
    float4* sh_c;         // light contribution xyz, -
    int32_t* hit_mesh;    // meshid of the first hit (bounce 0 only; feeds the AOVs)
    float4* aov_nd;       // AOV normal.xyz + depth, per PIXEL (PathTracingImplBase::aov_, pathtracing.h:96-98)
    float4* aov_am;       // AOV albedo.rgb + meshid, per pixel
};

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

// ---------------------------------------------------------------------------------------------
// generate / shade / gather
// ---------------------------------------------------------------------------------------------
#ifndef ATN_PIXEL_BLOCKS
#define ATN_PIXEL_BLOCKS 1
#endif
struct FrameParams {
    int32_t width, height;          // full image
    int32_t x0, y0, rw, rh;         // region rendered by this call; slot s <-> (x0 + s % rw, y0 + s / rw)
    uint32_t frame;                 // host frame index + sample index (pathtracing_impl.h:76-81)
    int32_t max_bounce, rr_bounce;
    uint32_t seed_count;
    int32_t batch;                  // camera samples traced as one wavefront: slot = k * (rw * rh) + pixel, sample k uses frame + k
    // tile-set mode (atn_render_tiles), tile > 0: the wavefront covers the tiles g = j * tile_world + tile_rank (j = 0 ..
    // rh - 1, scanline order over a grid of tiles_x tiles per row); rw = tile * tile slots per tile, rh = number of tiles.
    // Slots of an edge tile that fall outside the frame are never queued.
    int32_t tile, tiles_x, tile_world, tile_rank;
    // region mode: consecutive slots walk 8 x 4 pixel blocks instead of scanlines (a warp's 32 primary rays and first hits
    // are neighbours in both directions).  Only where nothing indexes path state by pixel: not with the bounce tap, not
    // in the AO / SVGF / ReSTIR renderers.
    int32_t swizzle;
};

// Where a path slot sits: pixel of the region, sample of the batch.  Samples are independent in the reference (one
// `render` pass per sample); a wavefront of one 1080p sample (2 M rays per launch) leaves ~6 % on the table against
// one of 8 M rays (launch tails, refill granularity), so several samples share a wavefront when the frame is small.
struct SlotInfo { int ix, iy, pix, k; uint32_t frame; bool valid; };
__device__ __forceinline__ SlotInfo slot_info(const FrameParams& fp, int slot)
{
    const int n_pix = fp.rw * fp.rh;
    SlotInfo si;
    si.k = (fp.batch > 1) ? slot / n_pix : 0;
    const int p = slot - si.k * n_pix;
    if (fp.tile > 0) {
        const int j = p / fp.rw, q = p - j * fp.rw;
        const int g = j * fp.tile_world + fp.tile_rank;
        si.ix = (g % fp.tiles_x) * fp.tile + q % fp.tile;
        si.iy = (g / fp.tiles_x) * fp.tile + q / fp.tile;
        si.valid = si.ix < fp.width && si.iy < fp.height;
    }
    else if (fp.swizzle) {
        const int blk = p >> 5, in = p & 31, per_row = fp.rw >> 3;
        const int by = blk / per_row, bx = blk - by * per_row;
        si.ix = fp.x0 + bx * 8 + (in & 7);
        si.iy = fp.y0 + by * 4 + (in >> 3);
        si.valid = true;
    }
    else {
        si.ix = fp.x0 + p % fp.rw;
        si.iy = fp.y0 + p / fp.rw;
        si.valid = true;
    }
    si.pix = si.iy * fp.width + si.ix;
    si.frame = fp.frame + (uint32_t)si.k;
    return si;
}

// Append `value` for the lanes with `pred` to a device queue: one atomicAdd per warp.
__device__ __forceinline__ int warp_append(bool pred, int32_t* counter)
{
    const unsigned mask = __ballot_sync(0xffffffffu, pred);
    if (mask == 0u) return -1;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(mask) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    return pred ? base + __popc(mask & ((1u << lane) - 1u)) : -1;
}

// GeneratePath (pathtracing_impl.h:65-110).  Region mode: path i of bounce 0 is slot i (no queue).  Tile-set mode: the
// slots inside the frame are appended to the bounce-0 queue.
__global__ void k_generate(FrameParams fp, AtnCameraParameter cam, const uint32_t* __restrict__ seeds, PathBuffers pb,
                           Counters* cnt)
{
    const int n = fp.rw * fp.rh * fp.batch;
    const int rounds = (n + blockDim.x * gridDim.x - 1) / (blockDim.x * gridDim.x);
    for (int it = 0; it < rounds; it++) {
        const int s = (it * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        bool live = false;
        if (s < n) {
            const SlotInfo si = slot_info(fp, s);
            live = si.valid;
            if (live) {
                const int ix = si.ix, iy = si.iy;
                const uint32_t idx = (uint32_t)si.pix;
                Sampler sp = sampler_for_path(si.frame, seeds[idx % fp.seed_count]);
                const float r1 = sp.next();
                const float r2 = sp.next();
                const float fs = (ix + r1) / (float)(cam.width);
                const float ft = (iy + r2) / (float)(cam.height);
                const Ray r = camera_sample(cam, fs, ft);
                pb.ray_o[s] = make_float4(r.org.x, r.org.y, r.org.z, 0.0f);
                pb.ray_d[s] = make_float4(r.dir.x, r.dir.y, r.dir.z, 0.0f);
                pb.throughput[s] = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
                pb.contrib[s] = make_float4(0.0f, 0.0f, 0.0f, 1.0f);
                pb.contrib_late[s] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                pb.sampler_dim[s] = sp.dim;
                pb.flags[s] = 0u;
                pb.pend[s] = 0;
            }
        }
        if (fp.tile > 0) {
            const int pos = warp_append(live, &cnt->n_active[0]);
            if (live) pb.queue[0][pos] = s;
        }
    }
    if (fp.tile == 0 && blockIdx.x == 0 && threadIdx.x == 0) cnt->n_active[0] = n;
}


#ifndef ATN_SURFACE_MIN_BLOCKS
#define ATN_SURFACE_MIN_BLOCKS 8
#endif
#ifndef ATN_NEE_MIN_BLOCKS
#define ATN_NEE_MIN_BLOCKS 8
#endif
#ifndef ATN_BSDF_MIN_BLOCKS
#define ATN_BSDF_MIN_BLOCKS 8
#endif
#ifndef ATN_SHADE_BLOCK
#define ATN_SHADE_BLOCK 128
#endif
constexpr int kShadeBlock = ATN_SHADE_BLOCK;
#ifndef ATN_WAVEFRONT_PATHS
#define ATN_WAVEFRONT_PATHS (32 * 1024 * 1024)    // paths per wavefront the sample batching aims at (8 M: -1.1 % on the bench frame)
#endif
#ifndef ATN_MAX_BATCH
#define ATN_MAX_BATCH 16
#endif

// Which paths a shade kernel works on.  The main pass of a bounce takes the bounce's trace queue and fills q_lit; the side
// pass takes the paths the exact walk has just decided (q_exact) and fills its own lit queue.  Both append to the same
// next-bounce and shadow queues.
struct ShadeQueues {
    const int32_t* q_in;  const int32_t* n_in;     // paths of the surface stage (q_in == nullptr: path i = slot i)
    int32_t* q_lit;       int32_t* n_lit;          // survivors of the surface stage = the paths of the NEE / BSDF stages
    int32_t skip_tag;                              // main pass next to a side pass: paths with pend == skip_tag are the side pass's
    int32_t* work;                                 // three zeroed work cursors: surface, NEE, BSDF stage (shade_chunk)
};

// Work distribution of the shade kernels: every warp takes 32 consecutive queue entries at a time from a cursor until
// the queue is empty.  A fixed share per block (the grid-stride loop these kernels had) makes a kernel as slow as its
// slowest block, and next to the side pass some blocks start late: k_side's blocks take register space that a full
// grid of these kernels needs, the blocks left without a slot started only when others had finished, and a side pass
// of 0.23 ms cost 0.15 ms per bounce of a 2 M-path wavefront (7 % of an SVGF or ReSTIR frame, 4 % of the bench frame).
// Measured on the bench frame (33 M-entry queues, one atomic per fetch on a single address): 32 / 64 / 128 / 256 / 384
// entries per fetch 1 963 / 1 961 / 1 953 / 1 934 / 1 922 Mrays/s -- the finer the better; sixteen cursors over sixteen
// segments of the queue instead of one: no gain.  A queue with no more than 32 entries per warp of the grid needs no
// cursor: warp w takes entries 32 w .. (a Cornell-box frame).  `*base` starts at -1.
__device__ __forceinline__ bool shade_chunk(int32_t* cursor, int count, int* base)
{
    int b;
    if (gridDim.x * blockDim.x >= count) {
        if (*base >= 0) return false;
        b = (blockIdx.x * blockDim.x + threadIdx.x) & ~31;
    }
    else {
        b = 0;
        if ((threadIdx.x & 31) == 0) b = atomicAdd(cursor, 32);
        b = __shfl_sync(0xffffffffu, b, 0);
    }
    *base = b;
    return b < count;
}

__device__ __forceinline__ Sampler path_sampler(const FrameParams& fp, const uint32_t* seeds, const PathBuffers& pb, int slot)
{
    const SlotInfo si = slot_info(fp, slot);
    Sampler sp = sampler_for_path(si.frame, seeds[(uint32_t)si.pix % fp.seed_count]);
    sp.dim = pb.sampler_dim[slot];
    return sp;
}

// The reference's shade kernel (pathtracing_impl.cu:18-238) is split in three so that each piece fits
// the instruction caches and runs only for the paths that need it ("Shade stage").
//
// Stage 1, all traced paths of the bounce: ShadeMiss for misses (pt_standard_impl.cu:330-362),
// hit evaluation + material + implicit light for hits.  Survivors go to q_lit with a Surface record.
// kSvgf: aten::SVGFRenderer::Shade instead of PathTracing::shade (see shade_surface); `w2c` is then the
// world-to-clip matrix of the frame and pb.aov_nd / pb.aov_am point at the current SVGF AOV set.
// The three stages of one path, as device functions: the wavefront kernels below run one stage over a queue, the side
// pass (k_side) runs all three for a path right after the exact walk has decided its hit.
template <bool kTap, bool kSvgf>
__device__ __forceinline__ bool surface_stage(const FrameParams& fp, const AtnCameraParameter& cam, const SceneView& sc, const PathBuffers& pb,
                                              int slot, int bounce, AtnBounceRecord* tap, int tap_cap, const AtnMat4& w2c)
{
    const float4 o = pb.ray_o[slot];
    const float4 d = pb.ray_d[slot];
    const float4 h0 = pb.hit0[slot];
    const int2 h1 = pb.hit1[slot];
    const float4 th = pb.throughput[slot];
    const float4 cb = pb.contrib[slot];
    const uint32_t fl = pb.flags[slot];
    Ray ray; ray.org = V3(o.x, o.y, o.z); ray.dir = V3(d.x, d.y, d.z);
    PathState st;
    st.throughput = V3(th.x, th.y, th.z); st.pdfb = th.w;
    st.contrib = V3(cb.x, cb.y, cb.z);
    st.terminated = false;
    st.singular = (fl & 2u) != 0u;
    st.last_specular = (fl & 4u) != 0u;
    // AOVs of the first hit are kept per pixel (the reference indexes them by pixel too)
    // (of the last sample of a batch: the buffers hold "the last sample rendered")
    const SlotInfo si = slot_info(fp, slot);
    const int pix = si.pix;
    const bool aov = bounce <= 1 && si.k == fp.batch - 1;
    AtnVec4* aov_nd = aov ? reinterpret_cast<AtnVec4*>(pb.aov_nd) + pix : nullptr;
    AtnVec4* aov_am = aov ? reinterpret_cast<AtnVec4*>(pb.aov_am) + pix : nullptr;
    bool lit = false;
    if (h1.x >= 0) {
        Surface sf;
        lit = shade_surface(sc, ray, h1.x, __float_as_int(h0.w), h0.y, h0.z, h0.x, h1.y, bounce, &st, &sf,
                            aov_nd, aov_am, bounce == 0 ? pb.hit_mesh[slot] : -1, kSvgf ? &w2c : nullptr);
        if (lit) {
            pb.surf0[slot] = make_float4(sf.p.x, sf.p.y, sf.p.z, sf.u);
            pb.surf1[slot] = make_float4(sf.n.x, sf.n.y, sf.n.z, sf.v);
            pb.surf2[slot] = make_float4(sf.albedo.x, sf.albedo.y, sf.albedo.z, __int_as_float(sf.mtrlid));
        }
    }
    else {
        shade_miss(sc, cam, si.ix, si.iy, fp.width, fp.height, bounce, ray, &st, aov_nd, aov_am);
    }
    if (!lit) {
        pb.contrib[slot] = make_float4(st.contrib.x, st.contrib.y, st.contrib.z, cb.w);
        pb.flags[slot] = 1u | (st.singular ? 2u : 0u);
    }
    if (kTap) {
        AtnBounceRecord& rec = tap[(size_t)slot * tap_cap + bounce];
        rec.ray.org = AtnVec3{ o.x, o.y, o.z };
        rec.ray.dir = AtnVec3{ d.x, d.y, d.z };
        rec.hit.t = h0.x; rec.hit.objid = h1.x; rec.hit.mtrlid = h1.y; rec.hit.meshid = -1;
        rec.hit.primid = __float_as_int(h0.w); rec.hit.a = h0.y; rec.hit.b = h0.z; rec.hit.isVoxel = 0;
        rec.throughput[0] = st.throughput.x; rec.throughput[1] = st.throughput.y; rec.throughput[2] = st.throughput.z;
        rec.pdfb = st.pdfb;
        rec.contrib[0] = st.contrib.x; rec.contrib[1] = st.contrib.y; rec.contrib[2] = st.contrib.z;
        rec.flags = 1 | (lit ? 0 : 2);
    }
    return lit;
}

__device__ __forceinline__ void nee_stage(const FrameParams& fp, const SceneView& sc, const uint32_t* __restrict__ seeds, const PathBuffers& pb, int slot)
{
    const float4 d = pb.ray_d[slot];
    const float4 s0 = pb.surf0[slot];
    const float4 s1 = pb.surf1[slot];
    const float4 s2 = pb.surf2[slot];
    const float4 th = pb.throughput[slot];
    Surface sf;
    sf.p = V3(s0.x, s0.y, s0.z); sf.u = s0.w; sf.n = V3(s1.x, s1.y, s1.z); sf.v = s1.w;
    sf.albedo = V3(s2.x, s2.y, s2.z); sf.mtrlid = __float_as_int(s2.w);
    PathState st;
    st.throughput = V3(th.x, th.y, th.z); st.pdfb = th.w;
    st.sampler = path_sampler(fp, seeds, pb, slot);
    ShadowRayOut sh;
    shade_nee(sc, V3(d.x, d.y, d.z), sf, &st, &sh);
    pb.sampler_dim[slot] = st.sampler.dim;
    // dist < 0 marks "no shadow ray"
    pb.sh_o[slot] = make_float4(sh.org.x, sh.org.y, sh.org.z, sh.active ? sh.dist : -1.0f);
    if (sh.active) pb.sh_d[slot] = make_float4(sh.dir.x, sh.dir.y, sh.dir.z, __int_as_float(sh.light_id));
    // ShadowRay::lightcontrib is reset whenever a light was sampled (FillShadowRay, pathtracing_impl.h:241) and stays
    // in the buffer: the ReSTIR visibility step of the next frame adds it again (restir_shadow_ray)
    // .w: the state of HitTestToTargetLight's loop for this shadow ray (ShadowPolicy<., kRetrace>): first lookup, whether the
    // surface asks for the stencil check, and the light's own object as the "last object hit"
    if (sh.sampled)
        pb.sh_c[slot] = make_float4(sh.lightcontrib.x, sh.lightcontrib.y, sh.lightcontrib.z,
                                    __int_as_float(shadow_state_pack(0, shadow_needs_stencil(sc, sf.mtrlid), shadow_light_obj(sc.lights[sh.light_id]))));
}

template <bool kTap>
__device__ __forceinline__ void bsdf_stage(const FrameParams& fp, const SceneView& sc, const uint32_t* __restrict__ seeds, const PathBuffers& pb,
                                           int slot, int bounce, AtnBounceRecord* tap, int tap_cap, bool* cont_out, bool* shadow_out)
{
    const float4 d = pb.ray_d[slot];
    const float4 s0 = pb.surf0[slot];
    const float4 s1 = pb.surf1[slot];
    const float4 s2 = pb.surf2[slot];
    const float4 th = pb.throughput[slot];
    const uint32_t fl = pb.flags[slot];
    Surface sf;
    sf.p = V3(s0.x, s0.y, s0.z); sf.u = s0.w; sf.n = V3(s1.x, s1.y, s1.z); sf.v = s1.w;
    sf.albedo = V3(s2.x, s2.y, s2.z); sf.mtrlid = __float_as_int(s2.w);
    PathState st;
    st.throughput = V3(th.x, th.y, th.z); st.pdfb = th.w;
    st.sampler = path_sampler(fp, seeds, pb, slot);
    st.terminated = false;
    st.singular = (fl & 2u) != 0u;
    Ray next; next.org = sf.p; next.dir = V3(d.x, d.y, d.z);
    shade_bsdf(sc, V3(d.x, d.y, d.z), sf, bounce, fp.rr_bounce, &st, &next);
    const bool cont = !st.terminated && (bounce + 1 < fp.max_bounce);
    const bool shadow = !st.terminated && pb.sh_o[slot].w >= 0.0f;
    pb.throughput[slot] = make_float4(st.throughput.x, st.throughput.y, st.throughput.z, st.pdfb);
    pb.sampler_dim[slot] = st.sampler.dim;
    pb.flags[slot] = (st.terminated ? 1u : 0u) | (st.singular ? 2u : 0u) | (st.last_specular ? 4u : 0u);
    if (cont) {
        pb.ray_o[slot] = make_float4(next.org.x, next.org.y, next.org.z, 0.0f);
        pb.ray_d[slot] = make_float4(next.dir.x, next.dir.y, next.dir.z, 0.0f);
    }
    if (kTap) {
        AtnBounceRecord& rec = tap[(size_t)slot * tap_cap + bounce];
        rec.shadow_active = shadow ? 1 : 0;
        rec.shadow_visible = 0;
        if (shadow) {
            const float4 so = pb.sh_o[slot]; const float4 sd = pb.sh_d[slot]; const float4 sc4 = pb.sh_c[slot];
            rec.shadow_org[0] = so.x; rec.shadow_org[1] = so.y; rec.shadow_org[2] = so.z; rec.shadow_dist = so.w;
            rec.shadow_dir[0] = sd.x; rec.shadow_dir[1] = sd.y; rec.shadow_dir[2] = sd.z;
            rec.light_contrib[0] = sc4.x; rec.light_contrib[1] = sc4.y; rec.light_contrib[2] = sc4.z;
        }
        rec.throughput[0] = st.throughput.x; rec.throughput[1] = st.throughput.y; rec.throughput[2] = st.throughput.z;
        rec.pdfb = st.pdfb;
        rec.flags = 1 | (st.terminated ? 2 : 0);
    }
    *cont_out = cont; *shadow_out = shadow;
}

template <bool kTap, bool kSvgf>
__global__ void __launch_bounds__(kShadeBlock, ATN_SURFACE_MIN_BLOCKS)
k_surface(FrameParams fp, AtnCameraParameter cam, SceneView sc, PathBuffers pb, ShadeQueues sq, int bounce,
          AtnBounceRecord* tap, int tap_cap, AtnMat4 w2c)
{
    const int count = *sq.n_in;
    const int32_t* __restrict__ q_in = sq.q_in;
    int base = -1;
    while (shade_chunk(sq.work + 0, count, &base)) {
        const int i = base + (threadIdx.x & 31);
        bool lit = false;
        int slot = 0;
        if (i < count) {
            slot = q_in ? q_in[i] : i;
            // a path whose hit comes from the exact walk is the side pass's (k_side)
            const bool pending = sq.skip_tag != 0 && pb.pend[slot] == sq.skip_tag;
            if (!pending) lit = surface_stage<kTap, kSvgf>(fp, cam, sc, pb, slot, bounce, tap, tap_cap, w2c);
        }
        const int pos = warp_append(lit, sq.n_lit);
        if (lit) sq.q_lit[pos] = slot;
    }
}

// Stage 2, lit paths: light pick + light sample + NEE evaluation (FillShadowRay).  The shadow-ray
// candidate is stored per slot; stage 3 decides whether it is traced.
__global__ void __launch_bounds__(kShadeBlock, ATN_NEE_MIN_BLOCKS)
k_nee(FrameParams fp, SceneView sc, const uint32_t* __restrict__ seeds, PathBuffers pb, ShadeQueues sq, int bounce)
{
    const int count = *sq.n_lit;
    // (Regrouping the 128 paths of a block by the light picked and the material, so that a warp runs one light-sampling
    // routine and one BSDF, was built and measured: bit-identical images, -1.3 % with the NEE stage regrouped and -15 %
    // with the BSDF stage -- warps of one kind finish at different times and wait for each other at the block's
    // barriers, and a warp's path-state loads spread over four times as many lines.  Not kept.)
    int base = -1;
    while (shade_chunk(sq.work + 1, count, &base)) {
        const int i = base + (threadIdx.x & 31);
        if (i < count) nee_stage(fp, sc, seeds, pb, sq.q_lit[i]);
    }
}

// Stage 3, lit paths: Russian roulette, BSDF sampling, next ray; queues the next bounce and the
// shadow ray (HitShadowRay drops the shadow ray of a path that terminated here).
template <bool kTap>
__global__ void __launch_bounds__(kShadeBlock, ATN_BSDF_MIN_BLOCKS)
k_bsdf(FrameParams fp, SceneView sc, const uint32_t* __restrict__ seeds, PathBuffers pb, Counters* cnt, ShadeQueues sq, int bounce,
       AtnBounceRecord* tap, int tap_cap)
{
    const int count = *sq.n_lit;
    int32_t* __restrict__ q_out = pb.queue[(bounce + 1) & 1];
    int base = -1;
    while (shade_chunk(sq.work + 2, count, &base)) {
        const int i = base + (threadIdx.x & 31);
        bool cont = false, shadow = false;
        int slot = 0;
        if (i < count) {
            slot = sq.q_lit[i];
            bsdf_stage<kTap>(fp, sc, seeds, pb, slot, bounce, tap, tap_cap, &cont, &shadow);
        }
        const int pos = warp_append(cont, &cnt->n_active[bounce + 1]);
        if (cont) q_out[pos] = slot;
        const int spos = warp_append(shadow, &cnt->n_shadow[bounce]);
        if (shadow) pb.q_shadow[spos] = slot;
    }
}

// The side pass of a bounce (launch_bounces): the paths whose closest hit depends on a near tie.  One thread per path: the
// reference's own walk decides the hit (ref_traverse), then the three shade stages run for that path right here, and the
// path joins the next bounce's queue and the shadow queue like any other.  One launch, so that it is done long before
// the main shade stage it runs next to.
template <bool kSvgf>
__global__ void __launch_bounds__(128)
k_side(RefView rv, ClosestPolicy pol, FrameParams fp, AtnCameraParameter cam, SceneView sc, const uint32_t* __restrict__ seeds,
       PathBuffers pb, Counters* cnt, int bounce, AtnMat4 w2c)
{
#ifdef ATN_EXP_EXACT_NOOP
    const int count = 0;
#else
    const int count = cnt->n_exact_closest[bounce];
#endif
    int32_t* __restrict__ q_out = pb.queue[(bounce + 1) & 1];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const int slot = pb.q_exact[i];
        V3 o, d;
        pol.world(slot, &o, &d);
        pol.finish_hit(slot, ref_traverse(rv, o, d, pol.t_max(slot)));
        if (!surface_stage<false, kSvgf>(fp, cam, sc, pb, slot, bounce, nullptr, 0, w2c)) continue;
        nee_stage(fp, sc, seeds, pb, slot);
        bool cont, shadow;
        bsdf_stage<false>(fp, sc, seeds, pb, slot, bounce, nullptr, 0, &cont, &shadow);
        if (cont) q_out[atomicAdd(&cnt->n_active[bounce + 1], 1)] = slot;
        if (shadow) pb.q_shadow[atomicAdd(&cnt->n_shadow[bounce], 1)] = slot;
    }
}

// ---------------------------------------------------------------------------------------------
// N1: ambient-occlusion renderer on the same path (idaten::AORenderer, src/libidaten/ao/ao.cu:20-121;
// host twin aten::AORenderer, src/libaten/renderer/ao/aorenderer.cpp).  It re-uses k_generate, the
// closest-hit traversal and k_gather; only the three small kernels below are its own.
// ---------------------------------------------------------------------------------------------
// After the primary trace: a miss stores ShadeByAOIfHitMiss() = 1 (ao.cu:27-33); a hit gets the AO
// sampler (aorenderer.cpp:27-29), its shading point and normal, and joins q_lit.
__global__ void __launch_bounds__(kShadeBlock)
k_ao_begin(FrameParams fp, SceneView sc, const uint32_t* __restrict__ seeds, PathBuffers pb, Counters* cnt)
{
    const int count = fp.rw * fp.rh * fp.batch;
    const int rounds = (count + blockDim.x * gridDim.x - 1) / (blockDim.x * gridDim.x);
    for (int it = 0; it < rounds; it++) {
        const int slot = (it * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        bool lit = false;
        if (slot < count) {
            const float4 h0 = pb.hit0[slot];
            const int2 h1 = pb.hit1[slot];
            if (h1.x >= 0) {
                const float4 o = pb.ray_o[slot];
                const float4 d = pb.ray_d[slot];
                Ray ray; ray.org = V3(o.x, o.y, o.z); ray.dir = V3(d.x, d.y, d.z);
                V3 p, n;
                ao_surface(sc, ray, h1.x, __float_as_int(h0.w), h0.y, h0.z, h0.x, h1.y, &p, &n);
                pb.surf0[slot] = make_float4(p.x, p.y, p.z, 0.0f);
                pb.surf1[slot] = make_float4(n.x, n.y, n.z, 0.0f);
                pb.contrib[slot] = make_float4(0.0f, 0.0f, 0.0f, 1.0f);
                pb.sampler_dim[slot] = 4 + 5 * 300;
                lit = true;
            }
            else {
                pb.contrib[slot] = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
            }
        }
        const int pos = warp_append(lit, &cnt->n_lit[0]);
        if (lit) pb.q_lit[pos] = slot;
    }
}

// One AO ray per lit path: Diffuse::sampleDirection + aten::ray(p, dir, n) (aorenderer_impl.h:70-71)
__global__ void __launch_bounds__(kShadeBlock)
k_ao_ray(FrameParams fp, const uint32_t* __restrict__ seeds, PathBuffers pb, Counters* cnt)
{
    const int count = cnt->n_lit[0];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const int slot = pb.q_lit[i];
        const SlotInfo si = slot_info(fp, slot);       // sample k of a batch is frame + k, as in the path tracer
        Sampler sp = ao_sampler_for_path(si.frame, seeds[(uint32_t)si.pix % fp.seed_count]);
        sp.dim = pb.sampler_dim[slot];
        const float4 s0 = pb.surf0[slot];
        const float4 s1 = pb.surf1[slot];
        V3 next_dir;
        const Ray ar = ao_ray(V3(s0.x, s0.y, s0.z), V3(s1.x, s1.y, s1.z), &sp, &next_dir);
        pb.sampler_dim[slot] = sp.dim;
        pb.ray_o[slot] = make_float4(ar.org.x, ar.org.y, ar.org.z, 0.0f);
        pb.ray_d[slot] = make_float4(ar.dir.x, ar.dir.y, ar.dir.z, 0.0f);
        pb.surf2[slot] = make_float4(next_dir.x, next_dir.y, next_dir.z, 0.0f);
    }
}

// Fold the AO ray's closest hit into the running value (aorenderer_impl.h:72-111).  A hit that is translucent by alpha
// restarts the ray behind it: its slot goes to `q_re` and the next pass traces it again (`last_lookup`: the 10th lookup).
__global__ void __launch_bounds__(kShadeBlock)
k_ao_accum(SceneView sc, PathBuffers pb, const int32_t* __restrict__ q_in, const int32_t* __restrict__ n_in,
           int32_t* __restrict__ q_re, int32_t* n_re, float radius, int last_lookup)
{
    const int count = *n_in;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const int slot = q_in[i];
        const float4 o = pb.ray_o[slot];
        const float4 d = pb.ray_d[slot];
        const float4 h0 = pb.hit0[slot];
        const int2 h1 = pb.hit1[slot];
        const float4 s1 = pb.surf1[slot];
        const float4 s2 = pb.surf2[slot];
        Ray ar; ar.org = V3(o.x, o.y, o.z); ar.dir = V3(d.x, d.y, d.z);
        float4 c = pb.contrib[slot];
        Ray next;
        const bool again = ao_accumulate(sc, &c.x, V3(s1.x, s1.y, s1.z), V3(s2.x, s2.y, s2.z), ar, radius,
                                         h1.x, __float_as_int(h0.w), h0.y, h0.z, h0.x, h1.y, last_lookup != 0, &next);
        pb.contrib[slot] = c;
        if (again) {
            pb.ray_o[slot] = make_float4(next.org.x, next.org.y, next.org.z, 0.0f);
            pb.ray_d[slot] = make_float4(next.dir.x, next.dir.y, next.dir.z, 0.0f);
            q_re[atomicAdd(n_re, 1)] = slot;
        }
    }
}

// `ao_color /= ao_num_rays`, spread over rgb (aorenderer_impl.h:114-116)
__global__ void __launch_bounds__(kShadeBlock)
k_ao_finish(PathBuffers pb, Counters* cnt, int num_rays)
{
    const int count = cnt->n_lit[0];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const int slot = pb.q_lit[i];
        float4 c = pb.contrib[slot];
        c.x /= num_rays; c.y = c.x; c.z = c.x;
        pb.contrib[slot] = c;
    }
}

// ---------------------------------------------------------------------------------------------
// N1: SVGF renderer on the same path (idaten::SVGFPathTracing, src/libidaten/svgf/svgf.cpp:28-160 and
// svgf_*.cu; host twin aten::SVGFRenderer, src/libaten/renderer/svgf/svgf.cpp).  The path-tracing half is
// the pipeline above with k_surface<.., true>; the image-space half is the five kernels below, one per
// pass of the reference (per-pixel functions in atn_svgf.cuh).  All of them are streaming stencils over
// W*H float4 planes: HBM / L2 bound.
// ---------------------------------------------------------------------------------------------
struct SvgfView {
    SvgfAov cur, prev;
    AtnVec4* atrous[2];
    AtnVec4* temporary;
    AtnVec4* scratch;             // variance estimation writes here (order-independent form), then becomes cur.color_variance
    const AtnVec4* motion_depth;  // or nullptr: static camera
};

// PrepareForDenoise; also folds the frame's queue counters into Totals (k_gather does that on the PT path)
__global__ void k_svgf_prepare(FrameParams fp, PathBuffers pb, SvgfView sv, int first_frame, const Counters* cnt, Totals* totals)
{
    const int n = fp.width * fp.height;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long c = 0, s = 0;
        unsigned long long e = 0;
        for (int b = 0; b < fp.max_bounce; b++) {
            c += (unsigned long long)cnt->n_active[b]; s += (unsigned long long)cnt->n_shadow[b];
            e += (unsigned long long)cnt->n_exact_closest[b] + (unsigned long long)cnt->n_exact_shadow[b];
        }
        totals->closest += c; totals->shadow += s; totals->exact += e;
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 c = pb.contrib[i]; const float4 l = pb.contrib_late[i];
        AtnVec4 cc; cc.x = c.x + l.x; cc.y = c.y + l.y; cc.z = c.z + l.z; cc.w = c.w;
        svgf_prepare(first_frame != 0, i, cc, sv.temporary, sv.cur);
    }
}

__global__ void k_svgf_temporal(FrameParams fp, PathBuffers pb, SvgfView sv)
{
    const int n = fp.width * fp.height;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 c = pb.contrib[i]; const float4 l = pb.contrib_late[i];
        AtnVec4 cc; cc.x = c.x + l.x; cc.y = c.y + l.y; cc.z = c.z + l.z; cc.w = c.w;
        svgf_temporal(i % fp.width, i / fp.width, fp.width, fp.height, 0.98f, 0.05f, cc, sv.cur, sv.prev, sv.motion_depth);
    }
}

__global__ void k_svgf_variance(FrameParams fp, SvgfView sv, float camera_distance)
{
    const int n = fp.width * fp.height;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        svgf_variance(i % fp.width, i / fp.width, fp.width, fp.height, camera_distance, sv.cur, sv.cur.color_variance, sv.scratch);
}

// ATN_SVGF_TEX: bit 0 = the normal-depth taps, bit 1 = the colour taps go through the texture pipe (a linear float4
// texture over all SVGF planes); 0 = everything through LDG.  The kernel is bound by the LSU pipe of L1 (84 128-bit
// loads per pixel); filter time per 1080p frame: 0: 1.83 ms, 1: 1.69, 2: 1.73, 3: 1.59 (default).
#ifndef ATN_SVGF_TEX
#define ATN_SVGF_TEX 3
#endif
struct SvgfTexPlane {
    cudaTextureObject_t tex; int32_t base;      // first texel of the plane inside the texture
    __device__ __forceinline__ AtnVec4 operator()(int32_t i) const
    {
        const float4 v = tex1Dfetch<float4>(tex, base + i);
        AtnVec4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w;
        return r;
    }
};
__global__ void k_svgf_atrous(FrameParams fp, SvgfView sv, float camera_distance, int iter, int iter_count, float4* image,
                              cudaTextureObject_t planes_tex, const AtnVec4* planes_base)
{
    const int n = fp.width * fp.height;
    // (selects, not sv.atrous[iter & 1]: a dynamic index would push the whole parameter struct to local memory)
    const AtnVec4* src = (iter == 0) ? sv.cur.color_variance : ((iter & 1) ? sv.atrous[1] : sv.atrous[0]);
    AtnVec4* next = (iter & 1) ? sv.atrous[0] : sv.atrous[1];
    const bool use_tex = planes_tex != 0;
    const SvgfTexPlane tnd{ planes_tex, (int32_t)(sv.cur.normal_depth - planes_base) };
    const SvgfTexPlane tcl{ planes_tex, (int32_t)(src - planes_base) };
    const SvgfPlane pnd{ sv.cur.normal_depth }, pcl{ src };
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        AtnVec4 film;
        bool done;
        const int ix = i % fp.width, iy = i / fp.width;
        if (use_tex && (ATN_SVGF_TEX & 3) == 1)
            done = svgf_atrous(iter, iter_count, ix, iy, fp.width, fp.height, camera_distance, sv.cur, src, next, sv.temporary, &film, tnd, pcl);
        else if (use_tex && (ATN_SVGF_TEX & 3) == 2)
            done = svgf_atrous(iter, iter_count, ix, iy, fp.width, fp.height, camera_distance, sv.cur, src, next, sv.temporary, &film, pnd, tcl);
        else if (use_tex && (ATN_SVGF_TEX & 3) == 3)
            done = svgf_atrous(iter, iter_count, ix, iy, fp.width, fp.height, camera_distance, sv.cur, src, next, sv.temporary, &film, tnd, tcl);
        else
            done = svgf_atrous(iter, iter_count, ix, iy, fp.width, fp.height, camera_distance, sv.cur, src, next, sv.temporary, &film, pnd, pcl);
        if (done) image[i] = make_float4(film.x, film.y, film.z, film.w);
    }
}

__global__ void k_svgf_copy(FrameParams fp, SvgfView sv)
{
    const int n = fp.width * fp.height;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) svgf_copy(i, sv.temporary, sv.cur);
}

// fills planes with aten::vec4's default value (0, 0, 0, 1): what the reference's buffers hold before frame 0
__global__ void k_fill_vec4(float4* p, size_t n, float4 v)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

// ---------------------------------------------------------------------------------------------
// N1: ReSTIR renderer on the same path (idaten::ReSTIRPathTracing, src/libidaten/restir/restir.cpp:47-123, restir.cu,
// restir_reuse.cu; per-pixel steps shared with aten::ReSTIRRenderer in restir_impl.h -> atn_restir.cuh).  Bounce 0 of a
// frame: k_restir_begin (hit evaluation, ReSTIRInfo, AOVs, implicit light), k_restir_candidate (one light candidate
// into the pixel's reservoir), k_bsdf (as on the path-tracing path), k_restir_visibility + the shadow traversal,
// k_restir_temporal (frames > 1), k_restir_spatial, k_restir_color.  Bounces > 0 are the path tracer's.
// ---------------------------------------------------------------------------------------------
struct RestirView {
    Reservoir* cur; Reservoir* other;           // other = previous frame's reservoirs, then the spatial destination
    RestirInfo* info_cur; RestirInfo* info_other;
    const AtnVec4* motion_depth;
};

__global__ void __launch_bounds__(kShadeBlock)
k_restir_begin(FrameParams fp, AtnCameraParameter cam, SceneView sc, PathBuffers pb, RestirView rv, Counters* cnt, AtnMat4 w2c)
{
    const int count = fp.width * fp.height;
    const int rounds = (count + blockDim.x * gridDim.x - 1) / (blockDim.x * gridDim.x);
    for (int it = 0; it < rounds; it++) {
        const int slot = (it * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        bool lit = false;
        if (slot < count) {
            const float4 o = pb.ray_o[slot];
            const float4 d = pb.ray_d[slot];
            const float4 h0 = pb.hit0[slot];
            const int2 h1 = pb.hit1[slot];
            const float4 th = pb.throughput[slot];
            const float4 cb = pb.contrib[slot];
            Ray ray; ray.org = V3(o.x, o.y, o.z); ray.dir = V3(d.x, d.y, d.z);
            PathState st;
            st.throughput = V3(th.x, th.y, th.z); st.pdfb = th.w;
            st.contrib = V3(cb.x, cb.y, cb.z);
            st.terminated = false; st.singular = false;
            AtnVec4* aov_nd = reinterpret_cast<AtnVec4*>(pb.aov_nd) + slot;
            AtnVec4* aov_am = reinterpret_cast<AtnVec4*>(pb.aov_am) + slot;
            rv.cur[slot].clear();                               // InitReSTIR (restir.cu:18-34)
            RestirInfo info = rv.info_cur[slot];
            info.clear();
            if (h1.x >= 0) {
                Surface sf;
                lit = restir_surface(sc, ray, h1.x, __float_as_int(h0.w), h0.y, h0.z, h0.x, h1.y, pb.hit_mesh[slot], 0, w2c, &st, &sf,
                                     &info, aov_nd, aov_am);
                if (lit) {
                    pb.surf0[slot] = make_float4(sf.p.x, sf.p.y, sf.p.z, sf.u);
                    pb.surf1[slot] = make_float4(sf.n.x, sf.n.y, sf.n.z, sf.v);
                    pb.surf2[slot] = make_float4(sf.albedo.x, sf.albedo.y, sf.albedo.z, __int_as_float(sf.mtrlid));
                }
            }
            else {
                shade_miss(sc, cam, slot % fp.width, slot / fp.width, fp.width, fp.height, 0, ray, &st, aov_nd, aov_am);
            }
            rv.info_cur[slot] = info;
            if (!lit) {
                pb.contrib[slot] = make_float4(st.contrib.x, st.contrib.y, st.contrib.z, cb.w);
                pb.flags[slot] = 1u;
            }
        }
        const int pos = warp_append(lit, &cnt->n_lit[0]);
        if (lit) pb.q_lit[pos] = slot;
    }
}

// GenerateInitialCandidate for the lit paths; no path-traced shadow ray at this bounce
__global__ void __launch_bounds__(kShadeBlock)
k_restir_candidate(FrameParams fp, SceneView sc, const uint32_t* __restrict__ seeds, PathBuffers pb, RestirView rv, Counters* cnt)
{
    const int count = cnt->n_lit[0];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const int slot = pb.q_lit[i];
        const float4 d = pb.ray_d[slot];
        const float4 s0 = pb.surf0[slot];
        const float4 s1 = pb.surf1[slot];
        const float4 s2 = pb.surf2[slot];
        const Mtrl mtrl = fill_material(sc, __float_as_int(s2.w));
        if (!(mtrl.is_singular() || mtrl.is_translucent())) {
            Sampler sp = path_sampler(fp, seeds, pb, slot);
            Reservoir res = rv.cur[slot];
            restir_initial_candidate(&res, mtrl, sc, V3(s0.x, s0.y, s0.z), V3(s1.x, s1.y, s1.z), V3(d.x, d.y, d.z), s0.w, s1.w, &sp);
            rv.cur[slot] = res;
            pb.sampler_dim[slot] = sp.dim;
        }
        float4 so = pb.sh_o[slot];
        so.w = -1.0f;
        pb.sh_o[slot] = so;
    }
}

// EvaluateVisibility, first half: the shadow ray of every live pixel whose reservoir holds a sample
__global__ void __launch_bounds__(kShadeBlock)
k_restir_visibility(FrameParams fp, PathBuffers pb, RestirView rv, Counters* cnt)
{
    const int count = fp.width * fp.height;
    const int rounds = (count + blockDim.x * gridDim.x - 1) / (blockDim.x * gridDim.x);
    for (int it = 0; it < rounds; it++) {
        const int slot = (it * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        bool shadow = false;
        if (slot < count && (pb.flags[slot] & 1u) == 0u) {
            Reservoir& res = rv.cur[slot];
            if (res.valid()) {
                RestirInfo info = rv.info_cur[slot];
                V3 org, dir; float dist;
                restir_shadow_ray(res, &info, &org, &dir, &dist);
                rv.info_cur[slot].point_to_light = info.point_to_light;
                pb.sh_o[slot] = make_float4(org.x, org.y, org.z, dist);
                pb.sh_d[slot] = make_float4(dir.x, dir.y, dir.z, __int_as_float(res.y));
                shadow = true;
            }
            else restir_visibility_result(&res, false);
        }
        const int spos = warp_append(shadow, &cnt->n_shadow[0]);
        if (shadow) pb.q_shadow[spos] = slot;
    }
}

// the shadow traversal of that step: ShadowPolicy's decision (HitShadowRay adds the buffered lightcontrib when the
// light is reached) + the second half of EvaluateVisibility
struct RestirShadowPolicy {
    static constexpr bool kNeedClass = true;
    ShadowPolicy<false> base;
    Reservoir* res;
    __device__ __forceinline__ void world(int slot, V3* o, V3* d) const { base.world(slot, o, d); }
    __device__ __forceinline__ float t_max(int slot) const { return base.t_max(slot); }
    __device__ __forceinline__ void load(int i, int* slot, V3* o, V3* d, float* t_max, int* mode, float* exit_dist) const
    {
        base.load(i, slot, o, d, t_max, mode, exit_dist);
    }
    __device__ __forceinline__ void finish_hit(int slot, const AtnHit& h) const
    {
        const float4 so = base.pb.sh_o[slot];
        const float4 sd = base.pb.sh_d[slot];
        const AtnLightParameter& light = base.sc.lights[__float_as_int(sd.w)];
        Ray sr; sr.org = V3(so.x, so.y, so.z); sr.dir = normalize(V3(sd.x, sd.y, sd.z));
        const bool vis = shadow_visible(base.sc, light, sr, so.w, h.objid, h.primid, h.a, h.b, h.t, h.mtrlid);
        if (vis) {
            const float4 c = base.pb.sh_c[slot];
            float4 cc = base.pb.contrib[slot];
            cc.x += c.x; cc.y += c.y; cc.z += c.z;
            base.pb.contrib[slot] = cc;
        }
        restir_visibility_result(&res[slot], vis);
    }
    __device__ __forceinline__ void finish_exact(int slot, const AtnHit& h) const { finish_hit(slot, h); }     // same stream, in order
    __device__ __forceinline__ bool finish(const BvhView& bv, int slot, const TraceResult& r, const Tie& tie) const
    {
        const AtnHit h = resolved(bv, r);
        if (tie.any && shadow_tie_matters(base.sc, base.sc.lights[__float_as_int(base.pb.sh_d[slot].w)], base.pb.sh_o[slot].w, h, tie)) return false;
        finish_hit(slot, h);
        return true;
    }
};

__global__ void __launch_bounds__(kShadeBlock)
k_restir_temporal(FrameParams fp, SceneView sc, const uint32_t* __restrict__ seeds, PathBuffers pb, RestirView rv)
{
    const int n = fp.width * fp.height;
    for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < n; slot += gridDim.x * blockDim.x) {
        if (pb.flags[slot] & 1u) continue;
        Sampler sp = path_sampler(fp, seeds, pb, slot);
        Reservoir res = rv.cur[slot];
        restir_temporal(slot, fp.width, fp.height, sc, &sp, &res, rv.info_cur[slot], rv.other, rv.info_other,
                        reinterpret_cast<const AtnVec4*>(pb.aov_am), rv.motion_depth);
        rv.cur[slot] = res;
        pb.sampler_dim[slot] = sp.dim;
    }
}

__global__ void __launch_bounds__(kShadeBlock)
k_restir_spatial(FrameParams fp, SceneView sc, const uint32_t* __restrict__ seeds, PathBuffers pb, RestirView rv)
{
    const int n = fp.width * fp.height;
    for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < n; slot += gridDim.x * blockDim.x) {
        if (pb.flags[slot] & 1u) continue;
        Sampler sp = path_sampler(fp, seeds, pb, slot);
        Reservoir res = rv.other[slot];
        restir_spatial(slot, fp.width, fp.height, sc, &sp, &res, rv.cur, rv.info_cur, reinterpret_cast<const AtnVec4*>(pb.aov_am));
        rv.other[slot] = res;
        pb.sampler_dim[slot] = sp.dim;
    }
}

__global__ void __launch_bounds__(kShadeBlock)
k_restir_color(FrameParams fp, SceneView sc, PathBuffers pb, RestirView rv)
{
    const int n = fp.width * fp.height;
    for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < n; slot += gridDim.x * blockDim.x) {
        if (pb.flags[slot] & 1u) continue;
        const float4 am = pb.aov_am[slot];
        AtnVec4 a; a.x = am.x; a.y = am.y; a.z = am.z; a.w = am.w;
        V3 c;
        if (restir_pixel_color(sc, rv.other[slot], rv.info_cur[slot], a, &c)) {
            float4 cc = pb.contrib[slot];
            cc.x += c.x; cc.y += c.y; cc.z += c.z;
            pb.contrib[slot] = cc;
        }
    }
}

// pt::gather (pathtracing_impl.cu:282-324) / FilmProgressive::put (film.cpp:61-71).  `mode` (AtnGatherMode):
//   OVERWRITE   the reference GPU renderer with progressive off: the pixel becomes the mean of THIS call's samples, w = 1
//               (`first`: the pixel's old value is ignored; `last`: the sum is divided by its count);
//   PROGRESSIVE FilmProgressive::put per sample: new = (n * old + sample) / (n + 1), w = n + 1;
//   SUM         xyz += sample, w += 1 (what the multi-GPU reduce combines).
__global__ void k_gather(FrameParams fp, PathBuffers pb, float4* image, int mode, int first, int last, const Counters* cnt, Totals* totals)
{
    const int n = fp.rw * fp.rh;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long c = 0, s = 0;
        unsigned long long e = 0;
        for (int b = 0; b < fp.max_bounce; b++) {
            c += (unsigned long long)cnt->n_active[b]; s += (unsigned long long)cnt->n_shadow[b];
            e += (unsigned long long)cnt->n_exact_closest[b] + (unsigned long long)cnt->n_exact_shadow[b];
        }
        totals->closest += c; totals->shadow += s; totals->exact += e;
    }
    for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < n; s += gridDim.x * blockDim.x) {
        FrameParams one = fp; one.batch = 1;
        const SlotInfo si = slot_info(one, s);
        if (!si.valid) continue;
        const int ix = si.ix, iy = si.iy;
        float4 cur = image[iy * fp.width + ix];
        if (mode == ATN_GATHER_OVERWRITE && first) cur = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        for (int k = 0; k < fp.batch; k++) {          // the samples of the batch, in sample order
            float4 c = pb.contrib[(size_t)k * n + s];
            { const float4 l = pb.contrib_late[(size_t)k * n + s]; c.x += l.x; c.y += l.y; c.z += l.z; }
            if (mode == ATN_GATHER_PROGRESSIVE) {
                const float nn = (float)((int32_t)cur.w);
                const float cx = c.x / 1.0f, cy = c.y / 1.0f, cz = c.z / 1.0f;   // col /= cnt, cnt = 1
                cur.x = (nn * cur.x + cx) / (nn + 1);
                cur.y = (nn * cur.y + cy) / (nn + 1);
                cur.z = (nn * cur.z + cz) / (nn + 1);
                cur.w = nn + 1;
            }
            else {
                cur.x += c.x; cur.y += c.y; cur.z += c.z; cur.w += 1.0f;
            }
        }
        if (mode == ATN_GATHER_OVERWRITE && last) {     // contrib /= sample; contrib.w = 1 (pathtracing_impl.cu:313-316)
            cur.x /= cur.w; cur.y /= cur.w; cur.z /= cur.w; cur.w = 1.0f;
        }
        image[iy * fp.width + ix] = cur;
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
template <class T>
struct DevBuf {
    T* p{ nullptr };
    size_t n{ 0 };
    ~DevBuf() { release(); }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
    cudaError_t resize(size_t count)
    {
        if (count <= n && p) return cudaSuccess;
        release();
        if (count == 0) return cudaSuccess;
        cudaError_t e = cudaMalloc(&p, count * sizeof(T));
        if (e == cudaSuccess) n = count;
        return e;
    }
    cudaError_t upload(const T* src, size_t count, cudaStream_t s)
    {
        cudaError_t e = resize(count);
        if (e != cudaSuccess || count == 0) return e;
        return cudaMemcpyAsync(p, src, count * sizeof(T), cudaMemcpyHostToDevice, s);
    }
};

struct AtnRenderer {
    int device{ 0 };
    cudaStream_t stream{ nullptr };
    int sm_count{ 148 };
    bool has_scene{ false };

    // scene
    DevBuf<AtnObjectParameter> objects;
    DevBuf<AtnMat4> matrices;
    DevBuf<AtnMaterialParameter> materials;
    DevBuf<AtnLightParameter> lights;
    int exp_bits{ 0 };
    DevBuf<float4> light_bounds;                 // SceneView::light_bounds (area lights: bounds of the light's object)
    std::vector<int32_t> h_light_obj;            // per light: the area light's object id, -1 otherwise
    std::vector<InstLeafW> h_insts;              // the top layer's leaves (object id + conservative world bounds)
    DevBuf<AtnTriangleParameter> triangles;
    DevBuf<AtnVec4> positions, normals;
    DevBuf<AtnVec4> texels;
    DevBuf<TexView> texviews;
    std::vector<TexView> h_texviews;           // host copy of the view table (atn_update_texture)
    DevBuf<QuadNode> nodes;
    DevBuf<TriLeaf> tris;
    DevBuf<InstLeafW> insts;
    // the reference's node lists as uploaded (exact walk of near ties): bottom lists first, list 0 behind them with room to grow
    DevBuf<AtnBvhNode> ref_nodes;
    DevBuf<uint32_t> ref_offsets;
    std::vector<uint32_t> h_ref_offsets;
    size_t ref_bottom_count{ 0 };      // nodes of lists 1.. (list 0 starts there)
    RefView rv{};
    int exact_ties{ 1 };               // 0: near ties are left as the front-to-back walk found them (atn_set_exact_ties)
    int exact_grid{ 148 * 4 };
    DevBuf<int32_t> q_exact, q_exact_s, q_exact_query, q_retrace0, q_retrace1;
    bool any_stencil_always{ false };
    bool alpha_surfaces{ false };      // a material can be translucent by alpha (baseColor.a < 1, or an albedo map with such a texel): AO rays loop
    std::vector<uint8_t> tex_has_alpha; // per texture: a texel with alpha < 1
    bool shadow_retrace{ false };      // alpha blending on or a stencil ALWAYS material in the scene: shadow rays run HitTestToTargetLight's loop
    // The exact walk is a latency-bound tail (a few thousand rays, hundreds of dependent steps each).  In the bounce loop it
    // runs on a second stream next to the main kernels instead of between them (launch_bounces).
    cudaStream_t side_stream{ nullptr };
    std::vector<cudaEvent_t> side_events;
    DevBuf<uint32_t> seeds;
    uint32_t seed_count{ 0 };
    SceneView sv{};
    BvhView bv{};
    AtnCameraParameter camera{};
    int32_t width{ 0 }, height{ 0 };
    float hit_distance_limit{ kInf };

    // host copies needed by atn_update_top_layer
    BottomInfo bottom;                  // roots / stack needs / root boxes of the encoded bottom trees (top-layer updates)
    size_t node_capacity{ 0 };          // QuadNodes the device array holds (bottom trees + room for the top layer)

    // path state
    DevBuf<float4> ray_o, ray_d, hit0, throughput, contrib, contrib_late, sh_o, sh_d, sh_c, image;
    DevBuf<int2> hit1;
    DevBuf<uint32_t> sampler_dim, flags;
    DevBuf<int32_t> hit_mesh;
    DevBuf<uint8_t> pend;
    DevBuf<float4> aov_nd, aov_am;
    DevBuf<int32_t> queue0, queue1, q_lit, q_shadow;
    DevBuf<float4> surf0, surf1, surf2;
    DevBuf<Counters> counters;
    DevBuf<Totals> totals;
    DevBuf<Visits> visits;
    bool stream_loads{ false };    // traversal variant for scenes beyond the L2 (k_trace<.., kStream = true>), set at upload
    // Render calls and device-buffer ray queries return without waiting for the GPU; their counters travel to this pinned
    // block behind the kernels and atn_get_stats collects them on demand (collect_stats).
    struct HostStats { Totals totals; Visits visits; Counters counters; };
    HostStats* h_stats{ nullptr };
    enum Pending { PEND_NONE, PEND_RENDER, PEND_AO, PEND_QUERY };
    Pending pending{ PEND_NONE };
    size_t pend_events{ 0 };       // profiling events recorded by the pending call
    uint64_t pend_ao_factor{ 0 };  // AO: rays per lit path of the call (num_rays * spp)
    uint64_t pend_ao_batch{ 1 };   // AO: samples in the call's last wavefront (n_lit[0] counts their lit paths)
    cudaEvent_t ev_done{ nullptr };
    PathBuffers pb{};

    // query scratch
    DevBuf<AtnRay> q_rays;
    DevBuf<float> q_f;
    DevBuf<AtnHit> q_hits;
    DevBuf<uint8_t> q_vis;

    // SVGF (N1): 12 planes of W*H float4 -- AOV sets 0 and 1 (normal_depth, albedo_meshid, color_variance, moment_tw),
    // a-trous ping / pong, temporary colour, variance scratch -- and the caller's motion vectors
    DevBuf<float4> svgf_planes, svgf_motion;
    int svgf_cv_plane[2]{ 2, 6 };      // which plane holds color_variance of each AOV set (swaps with the scratch plane)
    int svgf_scratch_plane{ 11 };
    bool svgf_has_motion{ false };
    DevBuf<Reservoir> restir_res[2];           // ReSTIR (N1): reservoirs and infos of the current / other frame
    DevBuf<RestirInfo> restir_info[2];
    int restir_cur{ 0 }, restir_target{ 0 };
    size_t restir_pixels{ 0 };
    cudaTextureObject_t svgf_tex{ 0 };  // linear float4 texture over svgf_planes (a-trous taps, ATN_SVGF_TEX)
    size_t svgf_pixels{ 0 };

    // tap
    AtnBounceRecord* tap_host{ nullptr };
    int tap_cap{ 0 };
    DevBuf<AtnBounceRecord> tap_dev;

    uint32_t frame{ 0 };
    int gather_mode{ ATN_GATHER_PROGRESSIVE };
    int profiling{ 0 };
    AtnStats stats{};
    cudaEvent_t ev0{ nullptr }, ev1{ nullptr };
    std::vector<cudaEvent_t> trace_events;
    int trace_grid{ 148 * 8 }, trace_grid_stream{ 148 * 6 };
    size_t l2_bytes{ 0 };
    int batch_cap{ 1 };                 // camera samples per wavefront (atn_upload_scene sizes the path buffers for it)
    size_t wavefront_paths{ (size_t)ATN_WAVEFRONT_PATHS };    // atn_set_wavefront_paths
    size_t path_capacity{ 0 };          // path slots allocated
    int stack_need{ 0 };                // worst-case traversal stack depth of the uploaded scene (WideBvh::max_stack)
    int surface_grid{ 148 * 8 }, nee_grid{ 148 * 8 }, bsdf_grid{ 148 * 8 };
};

// One traversal launch: the instantiation whose stack covers the scene's worst case (upload_bvh checked that one does),
// followed by the exact walk of the rays it could not decide (a few thousand of millions; nothing when exact ties are off).
// `exact_count` is a zeroed device counter of this launch, `exact_q` room for as many slots as the launch has rays.
template <class Policy>
static void launch_exact(AtnRenderer* r, cudaStream_t s, Policy pol, const int32_t* exact_q, const int32_t* exact_count)
{
    k_exact<Policy><<<r->exact_grid, 128, 0, s>>>(r->rv, pol, exact_q, exact_count);
}

template <class Policy>
static void launch_trace(AtnRenderer* r, cudaStream_t s, Policy pol, const int32_t* count_ptr, int count_imm, int32_t* fetch, Visits* visits,
                         int32_t* exact_q, int32_t* exact_count, bool exact_here = true)
{
    const bool small = r->stack_need <= kSmemStack + kLocalStackSmall;
    if (!r->exact_ties) { exact_q = nullptr; exact_count = nullptr; }
    const bool count = (r->profiling & 2) && visits;  // atn_set_profiling bit 2
#define ATN_LAUNCH_TRACE(LOCAL, STREAM, COUNT, GRID) \
    k_trace<Policy, LOCAL, STREAM, COUNT><<<GRID, kTraceBlock, 0, s>>>(r->bv, pol, count_ptr, count_imm, fetch, visits, exact_q, exact_count)
    if (r->stream_loads) {
        if (small) { if (count) ATN_LAUNCH_TRACE(kLocalStackSmall, true, true, r->trace_grid_stream); else ATN_LAUNCH_TRACE(kLocalStackSmall, true, false, r->trace_grid_stream); }
        else { if (count) ATN_LAUNCH_TRACE(kLocalStackLarge, true, true, r->trace_grid_stream); else ATN_LAUNCH_TRACE(kLocalStackLarge, true, false, r->trace_grid_stream); }
    }
    else {
        if (small) { if (count) ATN_LAUNCH_TRACE(kLocalStackSmall, false, true, r->trace_grid); else ATN_LAUNCH_TRACE(kLocalStackSmall, false, false, r->trace_grid); }
        else { if (count) ATN_LAUNCH_TRACE(kLocalStackLarge, false, true, r->trace_grid); else ATN_LAUNCH_TRACE(kLocalStackLarge, false, false, r->trace_grid); }
    }
#undef ATN_LAUNCH_TRACE
    if (exact_q && exact_here) launch_exact<Policy>(r, s, pol, exact_q, exact_count);
}

// (Re)binds what depends on the node array and the top layer: the linear node texture of the default kernel form, the
// instance leaves, the root and the stack bound.
// SceneView::light_bounds from the host copies of the lights' object ids and the top layer's leaves; called whenever either
// changes (upload, atn_update_top_layer, atn_update_lights).
static int refresh_light_bounds(AtnRenderer* r)
{
    const size_t n = r->h_light_obj.size();
    std::vector<float4> lb(2 * n, make_float4(0.0f, 0.0f, 0.0f, 0.0f));
    for (size_t l = 0; l < n; l++) {
        const int32_t obj = r->h_light_obj[l];
        if (obj < 0 || (r->exp_bits & 8)) continue;      // bit 8: no early exit for area-light shadow rays (same images, by construction)
        float mn[3] = { kInf, kInf, kInf }, mx[3] = { -kInf, -kInf, -kInf };
        bool found = false, finite = true;
        for (const InstLeafW& w : r->h_insts) {
            if (w.objid != obj) continue;
            found = true;
            for (int a = 0; a < 3; a++) {
                mn[a] = std::min(mn[a], w.wmin[a]); mx[a] = std::max(mx[a], w.wmax[a]);
                finite = finite && std::isfinite(w.wmin[a]) && std::isfinite(w.wmax[a]);
            }
        }
        if (!found || !finite) continue;
        lb[2 * l] = make_float4(mn[0], mn[1], mn[2], 1.0f);
        lb[2 * l + 1] = make_float4(mx[0], mx[1], mx[2], 1.0f);
    }
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    CUDA_TRY(r->light_bounds.upload(lb.data(), lb.size(), r->stream));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    r->sv.light_bounds = n ? reinterpret_cast<const AtnVec4*>(r->light_bounds.p) : nullptr;
    return ATN_OK;
}
static void remember_lights(AtnRenderer* r, const AtnLightParameter* l, uint32_t count)
{
    r->h_light_obj.assign(count, -1);
    for (uint32_t i = 0; i < count; i++) r->h_light_obj[i] = atn::shadow_light_obj(l[i]);
}

static int bind_top_layer(AtnRenderer* r, size_t node_count, const std::vector<InstLeafW>& insts, int32_t top_root, int32_t max_stack)
{
    if (max_stack > kSmemStack + kLocalStackLarge)
        return set_error(ATN_ERR_TOO_LARGE, "BVH needs a traversal stack of %d entries (limit %d)", max_stack, kSmemStack + kLocalStackLarge);
    r->stack_need = max_stack;
    r->h_insts = insts;
    if (int rc = refresh_light_bounds(r)) return rc;
    CUDA_TRY(r->insts.upload(insts.data(), insts.size(), r->stream));
    CUDA_TRY(cudaStreamSynchronize(r->stream));   // the caller's host arrays go out of scope
    r->bv.nodes = reinterpret_cast<const float4*>(r->nodes.p);
    if (r->bv.node_tex) { cudaDestroyTextureObject(r->bv.node_tex); r->bv.node_tex = 0; }
    // linear float4 texture over the node array; tex1Dfetch addresses 2^27 texels = 16 M QuadNodes (2 GB).  Only the
    // default form of the kernel reads nodes through it: a larger set (which is far beyond the L2 anyway) is served by
    // the streaming form, with plain loads.
    if (node_count * 8 > ((size_t)1 << 27)) r->stream_loads = true;
    if (ATN_NODE_TEX && node_count > 0 && !r->stream_loads) {
        cudaResourceDesc rd{}; rd.resType = cudaResourceTypeLinear; rd.res.linear.devPtr = r->nodes.p;
        rd.res.linear.desc = cudaCreateChannelDesc<float4>(); rd.res.linear.sizeInBytes = node_count * sizeof(QuadNode);
        cudaTextureDesc td{}; td.readMode = cudaReadModeElementType;
        CUDA_TRY(cudaCreateTextureObject(&r->bv.node_tex, &rd, &td, nullptr));
    }
    r->bv.tris = reinterpret_cast<const float4*>(r->tris.p);
    r->bv.insts = reinterpret_cast<const float4*>(r->insts.p);
    r->bv.matrices = r->matrices.p;
    r->bv.top_root = top_root;
    return ATN_OK;
}

// The reference's node lists as they are, for k_exact: bottom lists (1..) first, list 0 behind them with room to grow
// (a top-layer update replaces list 0 only).
static int bind_ref_lists(AtnRenderer* r)
{
    RefView& rv = r->rv;
    rv.nodes = r->ref_nodes.p; rv.offsets = r->ref_offsets.p;
    rv.objects = r->objects.p; rv.matrices = r->matrices.p; rv.triangles = r->triangles.p; rv.positions = r->positions.p;
    rv.t_min = r->bv.t_min;
    return ATN_OK;
}

static int upload_ref_top(AtnRenderer* r, const AtnBvhNode* top, uint32_t count)
{
    const size_t need = r->ref_bottom_count + count;
    if (need > r->ref_nodes.n) {
        DevBuf<AtnBvhNode> grown;
        CUDA_TRY(grown.resize(need + count + 64));
        if (r->ref_bottom_count)
            CUDA_TRY(cudaMemcpyAsync(grown.p, r->ref_nodes.p, r->ref_bottom_count * sizeof(AtnBvhNode), cudaMemcpyDeviceToDevice, r->stream));
        CUDA_TRY(cudaStreamSynchronize(r->stream));
        std::swap(grown.p, r->ref_nodes.p); std::swap(grown.n, r->ref_nodes.n);
    }
    if (count) CUDA_TRY(cudaMemcpyAsync(r->ref_nodes.p + r->ref_bottom_count, top, count * sizeof(AtnBvhNode), cudaMemcpyHostToDevice, r->stream));
    return bind_ref_lists(r);
}

static int upload_ref_lists(AtnRenderer* r, const AtnSceneBlob* scene)
{
    const uint32_t nl = scene->node_list_count;
    r->h_ref_offsets.assign(nl, 0u);
    size_t total = 0;
    for (uint32_t k = 1; k < nl; k++) { r->h_ref_offsets[k] = (uint32_t)total; total += scene->node_list_sizes[k]; }
    if (total + scene->node_list_sizes[0] >= ((size_t)1 << 32)) return set_error(ATN_ERR_TOO_LARGE, "node lists: more than 2^32 nodes");
    r->h_ref_offsets[0] = (uint32_t)total;
    r->ref_bottom_count = total;
    r->ref_nodes.release();
    CUDA_TRY(r->ref_nodes.resize(total + 2 * (size_t)scene->node_list_sizes[0] + 64));
    for (uint32_t k = 1; k < nl; k++)
        if (scene->node_list_sizes[k])
            CUDA_TRY(cudaMemcpyAsync(r->ref_nodes.p + r->h_ref_offsets[k], scene->node_lists[k], (size_t)scene->node_list_sizes[k] * sizeof(AtnBvhNode),
                                     cudaMemcpyHostToDevice, r->stream));
    CUDA_TRY(r->ref_offsets.upload(r->h_ref_offsets.data(), nl, r->stream));
    return upload_ref_top(r, scene->node_lists[0], scene->node_list_sizes[0]);
}

// `cache_path` (optional): file of the encoded bottom trees (atn_cache.h).  A file that holds the encoding of exactly these
// node lists, triangles and positions is used instead of encoding them again; otherwise they are encoded and the file is
// (re)written.  *cache_hit (optional) says which happened.
static int upload_bvh(AtnRenderer* r, const AtnSceneBlob* scene, const char* cache_path = nullptr, int* cache_hit = nullptr)
{
    {
        const int rc = upload_ref_lists(r, scene);
        if (rc != ATN_OK) return rc;
    }
    if (cache_hit) *cache_hit = 0;
    WideBvh enc;
    if (cache_path && cache_path[0]) {
        const uint64_t hash = source_hash(scene);
        CachedBottom cached;
        const std::string why = read_encoded(cache_path, hash, scene->node_list_count, &cached);
        if (why.empty()) {
            enc = encode_wide(scene, &r->bottom, &cached);
            if (cache_hit) *cache_hit = 1;
        }
        else {
            enc = encode_wide(scene, &r->bottom);
            if (enc.error.empty()) {
                const size_t nb = (size_t)r->bottom.node_count;
                const std::string werr = write_encoded(cache_path, hash, scene->node_list_count, r->bottom, enc.nodes.data(), enc.stack_need.data(), nb,
                                                       enc.tris.data(), enc.tris.size(), enc.slots_used, enc.replaced, enc.not_nested);
                if (!werr.empty()) return set_error(ATN_ERR_INVALID_ARGUMENT, "encoded-tree cache: %s", werr.c_str());
            }
        }
    }
    else enc = encode_wide(scene, &r->bottom);
    if (!enc.error.empty()) return set_error(ATN_ERR_INVALID_ARGUMENT, "node lists: %s", enc.error.c_str());
    {
        // nodes + triangles far beyond the L2: the traversal waits for device memory, not for issue slots -- use the variant
        // that issues an iteration's loads together (k_trace<.., kStream>).  ATEN_B200_STREAM=0/1 overrides (experiments).
        // Where the forms cross, measured with 20 M incoherent rays into n distinct 100 k-triangle meshes (13.4 MB each;
        // closest-hit Mrays/s, default / streaming): n = 14 (187 MB) 1 744 / 1 537, 25 (334 MB) 1 630 / 1 455,
        // 40 (535 MB) 1 301 / 1 339, 60 (802 MB) 859 / 1 394, 100 (1.3 GB) 722 / 1 266: the hot upper levels of the trees
        // stay in the 126 MB L2 long after the whole set has outgrown it, so the switch is at four times its size.
        const size_t set_bytes = enc.nodes.size() * sizeof(QuadNode) + enc.tris.size() * sizeof(TriLeaf);
        r->stream_loads = r->l2_bytes > 0 && set_bytes > 4 * r->l2_bytes;
        if (const char* ov = getenv("ATEN_B200_STREAM")) r->stream_loads = (ov[0] == '1');
    }
    // room for a top layer of twice today's size: an update with more instances grows the array (atn_update_top_layer)
    const size_t top_nodes = enc.nodes.size() - (size_t)r->bottom.node_count;
    r->nodes.release();
    CUDA_TRY(r->nodes.resize(enc.nodes.size() + top_nodes + 64));
    r->node_capacity = r->nodes.n;
    CUDA_TRY(cudaMemcpyAsync(r->nodes.p, enc.nodes.data(), enc.nodes.size() * sizeof(QuadNode), cudaMemcpyHostToDevice, r->stream));
    CUDA_TRY(r->tris.upload(enc.tris.data(), enc.tris.size(), r->stream));
    return bind_top_layer(r, enc.nodes.size(), enc.insts, enc.top_root, enc.max_stack);
}

// Shadow rays run HitTestToTargetLight's loop when alpha blending is on or a surface can ask for the stencil check.
static void update_shadow_retrace(AtnRenderer* r, const AtnMaterialParameter* m, uint32_t count)
{
    if (m) {
        r->any_stencil_always = false;
        for (uint32_t i = 0; i < count; i++) if (m[i].stencil_type == kStencilAlways) r->any_stencil_always = true;
    }
    r->shadow_retrace = r->sv.alpha_blending != 0 || r->any_stencil_always;
    if (m) {
        r->alpha_surfaces = false;
        for (uint32_t i = 0; i < count; i++) {
            const bool tex = m[i].albedoMap >= 0 && (size_t)m[i].albedoMap < r->tex_has_alpha.size() && r->tex_has_alpha[m[i].albedoMap];
            if (m[i].baseColor.w < 1.0f || tex) r->alpha_surfaces = true;
        }
    }
}

static bool texels_have_alpha(const AtnVec4* t, size_t n)
{
    for (size_t i = 0; i < n; i++) if (t[i].w < 1.0f) return true;
    return false;
}

static float scene_radius_of(const AtnSceneConfig& cfg)
{
    // ImageBasedLight::sample (ibl.h:105-110): ComputeDistanceToCoverBoundingSphere(Deg2Rad(30))
    const float* mn = cfg.scene_bbox_min; const float* mx = cfg.scene_bbox_max;
    const bool valid = !((mn[0] >= mx[0]) || (mn[1] >= mx[1]) || (mn[2] >= mx[2]));
    if (!valid) return 10000.0f;
    const V3 vmin(mn[0], mn[1], mn[2]), vmax(mx[0], mx[1], mx[2]);
    const V3 center = (vmin + vmax) * 0.5f;
    const float rad = length(vmax - center);
    const float theta = kPi * 30.0f / 180.0f;
    return rad / tanf(theta / 2);
}

static int alloc_paths(AtnRenderer* r, size_t n, size_t pixels)
{
    CUDA_TRY(r->ray_o.resize(n)); CUDA_TRY(r->ray_d.resize(n)); CUDA_TRY(r->hit0.resize(n)); CUDA_TRY(r->hit1.resize(n));
    CUDA_TRY(r->throughput.resize(n)); CUDA_TRY(r->contrib.resize(n)); CUDA_TRY(r->contrib_late.resize(n)); CUDA_TRY(r->sampler_dim.resize(n));
    CUDA_TRY(r->flags.resize(n)); CUDA_TRY(r->queue0.resize(n)); CUDA_TRY(r->queue1.resize(n));
    CUDA_TRY(r->sh_o.resize(n)); CUDA_TRY(r->sh_d.resize(n)); CUDA_TRY(r->sh_c.resize(n));
    CUDA_TRY(cudaMemsetAsync(r->sh_c.p, 0, n * sizeof(float4), r->stream));      // ShadowRay::lightcontrib starts at zero
    CUDA_TRY(r->q_lit.resize(n)); CUDA_TRY(r->q_shadow.resize(n)); CUDA_TRY(r->q_exact.resize(n));
    CUDA_TRY(r->q_exact_s.resize(n)); CUDA_TRY(r->pend.resize(n));
    CUDA_TRY(r->q_retrace0.resize(n)); CUDA_TRY(r->q_retrace1.resize(n));
    CUDA_TRY(r->hit_mesh.resize(n)); CUDA_TRY(r->aov_nd.resize(pixels)); CUDA_TRY(r->aov_am.resize(pixels));    // AOVs are per pixel
    CUDA_TRY(r->surf0.resize(n)); CUDA_TRY(r->surf1.resize(n)); CUDA_TRY(r->surf2.resize(n));
    CUDA_TRY(r->counters.resize(1));
    CUDA_TRY(r->totals.resize(1));
    CUDA_TRY(r->visits.resize(1));
    PathBuffers& pb = r->pb;
    pb.ray_o = r->ray_o.p; pb.ray_d = r->ray_d.p; pb.hit0 = r->hit0.p; pb.hit1 = r->hit1.p;
    pb.throughput = r->throughput.p; pb.contrib = r->contrib.p; pb.contrib_late = r->contrib_late.p; pb.sampler_dim = r->sampler_dim.p; pb.flags = r->flags.p;
    pb.queue[0] = r->queue0.p; pb.queue[1] = r->queue1.p;
    pb.sh_o = r->sh_o.p; pb.sh_d = r->sh_d.p; pb.sh_c = r->sh_c.p;
    pb.q_lit = r->q_lit.p; pb.q_shadow = r->q_shadow.p; pb.q_exact = r->q_exact.p; pb.q_exact_s = r->q_exact_s.p; pb.pend = r->pend.p; pb.q_retrace[0] = r->q_retrace0.p; pb.q_retrace[1] = r->q_retrace1.p;
    pb.hit_mesh = r->hit_mesh.p; pb.aov_nd = r->aov_nd.p; pb.aov_am = r->aov_am.p;
    pb.surf0 = r->surf0.p; pb.surf1 = r->surf1.p; pb.surf2 = r->surf2.p;
    return ATN_OK;
}

extern "C" ATN_API int atn_destroy(AtnRenderer* r);

// Everything atn_create sets up after the allocation of the handle; on failure atn_create destroys the half-built handle.
static int init_renderer(AtnRenderer* r)
{
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, r->device));
    r->sm_count = prop.multiProcessorCount;
    CUDA_TRY(cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking));
    {
        int lo = 0, hi = 0;
        CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CUDA_TRY(cudaStreamCreateWithPriority(&r->side_stream, cudaStreamNonBlocking, hi));     // its few blocks go first when an SM frees up
    }
    CUDA_TRY(cudaEventCreate(&r->ev0));
    CUDA_TRY(cudaEventCreate(&r->ev1));
    CUDA_TRY(cudaEventCreateWithFlags(&r->ev_done, cudaEventDisableTiming));
    CUDA_TRY(cudaMallocHost(&r->h_stats, sizeof(AtnRenderer::HostStats)));
    memset(r->h_stats, 0, sizeof(AtnRenderer::HostStats));
    int bps = 0;
    r->l2_bytes = (size_t)prop.l2CacheSize;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_trace<ClosestPolicy, kLocalStackSmall, false>, kTraceBlock, 0));
    r->trace_grid = r->sm_count * std::max(1, bps);
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_trace<ClosestPolicy, kLocalStackSmall, true>, kTraceBlock, 0));
    r->trace_grid_stream = r->sm_count * std::max(1, bps);
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_surface<false, false>, kShadeBlock, 0));
    r->surface_grid = r->sm_count * std::max(1, bps);
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_nee, kShadeBlock, 0));
    r->nee_grid = r->sm_count * std::max(1, bps);
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_bsdf<false>, kShadeBlock, 0));
    r->bsdf_grid = r->sm_count * std::max(1, bps);
    r->exact_grid = r->sm_count * 4;
    // experiments (two renderers side by side on one GPU): blocks per SM of the traversal and of the shade kernels
    if (const char* ov = getenv("ATEN_B200_TRACE_BPS")) r->trace_grid = r->sm_count * std::max(1, atoi(ov));
    if (const char* ov = getenv("ATEN_B200_SHADE_BPS")) r->surface_grid = r->nee_grid = r->bsdf_grid = r->sm_count * std::max(1, atoi(ov));
    if (const char* ov = getenv("ATEN_B200_EXACT_TIES")) r->exact_ties = (ov[0] != '0');     // experiments: cost of the exact walk
    if (const char* ov = getenv("ATEN_B200_EXP")) r->exp_bits = atoi(ov);                     // experiments, none changes a result: 1 = the exact walk in line on the main stream, 8 = no area-light early exit
    return ATN_OK;
}

// End of an asynchronous call: the device counters follow the kernels into pinned host memory; nothing waits here.
static int end_call(AtnRenderer* r, AtnRenderer::Pending kind, uint64_t launches, uint64_t samples, size_t n_events)
{
    cudaStream_t s = r->stream;
    CUDA_TRY(cudaEventRecord(r->ev1, s));
    CUDA_TRY(cudaGetLastError());
    if (kind != AtnRenderer::PEND_QUERY) CUDA_TRY(cudaMemcpyAsync(&r->h_stats->totals, r->totals.p, sizeof(Totals), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaMemcpyAsync(&r->h_stats->visits, r->visits.p, sizeof(Visits), cudaMemcpyDeviceToHost, s));
    if (kind != AtnRenderer::PEND_RENDER) CUDA_TRY(cudaMemcpyAsync(&r->h_stats->counters, r->counters.p, sizeof(Counters), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaEventRecord(r->ev_done, s));
    r->stats.kernel_launches = launches;
    r->stats.camera_samples = samples;
    r->pend_events = n_events;
    r->pending = kind;
    return ATN_OK;
}

// Waits for the pending call (if any) and fills r->stats from what it left in the pinned block.
static int collect_stats(AtnRenderer* r)
{
    if (r->pending == AtnRenderer::PEND_NONE) return ATN_OK;
    CUDA_TRY(cudaSetDevice(r->device));
    CUDA_TRY(cudaEventSynchronize(r->ev_done));
    const AtnRenderer::Pending kind = r->pending;
    r->pending = AtnRenderer::PEND_NONE;
    float ms = 0.0f;
    CUDA_TRY(cudaEventElapsedTime(&ms, r->ev0, r->ev1));
    r->stats.device_ms = ms;
    r->stats.nodes_visited = r->h_stats->visits.nodes;
    r->stats.tris_tested = r->h_stats->visits.tris;
    if (kind == AtnRenderer::PEND_QUERY) { r->stats.trace_ms = ms; r->stats.exact_rays = (uint64_t)r->h_stats->counters.n_exact_query; return ATN_OK; }
    r->stats.closest_rays = r->h_stats->totals.closest;
    r->stats.shadow_rays = r->h_stats->totals.shadow;
    r->stats.exact_rays = r->h_stats->totals.exact;
    if (kind == AtnRenderer::PEND_AO)      // AO rays = lit paths per sample (from the last wavefront) x rays per path x samples
        r->stats.shadow_rays = (uint64_t)r->h_stats->counters.n_lit[0] * r->pend_ao_factor / r->pend_ao_batch;
    r->stats.trace_ms = 0.0f;
    float tot = 0.0f;
    for (size_t i = 0; i + 3 < r->pend_events; i += 4) {
        float a = 0.0f, b2 = 0.0f;
        cudaEventElapsedTime(&a, r->trace_events[i], r->trace_events[i + 1]);
        cudaEventElapsedTime(&b2, r->trace_events[i + 2], r->trace_events[i + 3]);
        tot += a + b2;
    }
    r->stats.trace_ms = tot;
    return ATN_OK;
}

extern "C" {

ATN_API const char* atn_last_error(void) { return g_last_error.c_str(); }
ATN_API const char* atn_version(void) { return "aten_b200 0.1 (sm_100a)"; }

ATN_API int atn_create(int device_ordinal, AtnRenderer** out)
{
    if (!out) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_create: out is NULL");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return set_error(ATN_ERR_CUDA, "atn_create: no CUDA device (%s); this library has no CPU path",
                         e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device_ordinal < 0 || device_ordinal >= ndev)
        return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_create: device %d out of range [0,%d)", device_ordinal, ndev);
    CUDA_TRY(cudaSetDevice(device_ordinal));
    AtnRenderer* r = new AtnRenderer();
    r->device = device_ordinal;
    const int rc = init_renderer(r);
    if (rc != ATN_OK) { const std::string msg = g_last_error; atn_destroy(r); g_last_error = msg; return rc; }   // *out untouched
    *out = r;
    return ATN_OK;
}

ATN_API int atn_destroy(AtnRenderer* r)
{
    if (!r) return ATN_OK;
    cudaSetDevice(r->device);
    if (r->stream) cudaStreamSynchronize(r->stream);
    for (auto ev : r->trace_events) cudaEventDestroy(ev);
    if (r->ev0) cudaEventDestroy(r->ev0);
    if (r->ev1) cudaEventDestroy(r->ev1);
    if (r->ev_done) cudaEventDestroy(r->ev_done);
    for (auto ev : r->side_events) cudaEventDestroy(ev);
    if (r->side_stream) { cudaStreamSynchronize(r->side_stream); cudaStreamDestroy(r->side_stream); }
    if (r->h_stats) cudaFreeHost(r->h_stats);
    if (r->bv.node_tex) cudaDestroyTextureObject(r->bv.node_tex);
    if (r->svgf_tex) cudaDestroyTextureObject(r->svgf_tex);
    cudaStream_t s = r->stream;
    delete r;
    if (s) cudaStreamDestroy(s);
    return ATN_OK;
}

static int upload_scene_impl(AtnRenderer* r, const AtnSceneBlob* scene, const AtnCameraParameter* camera,
                             int32_t width, int32_t height, const char* cache_path, int* cache_hit)
{
    if (!r || !scene || !camera) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_upload_scene: NULL argument");
    if (width <= 0 || height <= 0) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_upload_scene: bad size %dx%d", width, height);
    if ((int64_t)width * height > (int64_t)1 << 28) return set_error(ATN_ERR_TOO_LARGE, "atn_upload_scene: image too large");
    if (scene->node_list_count == 0 || !scene->node_lists) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_upload_scene: no node lists");
    CUDA_TRY(cudaSetDevice(r->device));
    cudaStream_t s = r->stream;
    CUDA_TRY(r->objects.upload(scene->objects, scene->object_count, s));
    CUDA_TRY(r->matrices.upload(scene->matrices, scene->matrix_count, s));
    CUDA_TRY(r->materials.upload(scene->materials, scene->material_count, s));
    CUDA_TRY(r->lights.upload(scene->lights, scene->light_count, s));
    remember_lights(r, scene->lights, scene->light_count);
    CUDA_TRY(r->triangles.upload(scene->triangles, scene->triangle_count, s));
    CUDA_TRY(r->positions.upload(scene->positions, scene->vertex_count, s));
    CUDA_TRY(r->normals.upload(scene->normals, scene->vertex_count, s));
    // textures: one texel pool + a table of views
    size_t total = 0;
    for (uint32_t i = 0; i < scene->texture_count; i++) total += (size_t)scene->textures[i].width * scene->textures[i].height;
    CUDA_TRY(r->texels.resize(total));
    std::vector<TexView> tv(scene->texture_count);
    size_t off = 0;
    for (uint32_t i = 0; i < scene->texture_count; i++) {
        const AtnTextureDesc& t = scene->textures[i];
        const size_t n = (size_t)t.width * t.height;
        tv[i].texels = t.texels ? r->texels.p + off : nullptr;
        tv[i].w = t.width; tv[i].h = t.height; tv[i].filter = t.filter_mode; tv[i].address = t.address_mode;
        if (t.texels && n) CUDA_TRY(cudaMemcpyAsync(r->texels.p + off, t.texels, n * sizeof(AtnVec4), cudaMemcpyHostToDevice, s));
        off += n;
    }
    r->tex_has_alpha.assign(scene->texture_count, 0);
    for (uint32_t i = 0; i < scene->texture_count; i++) {
        const AtnTextureDesc& t = scene->textures[i];
        if (t.texels) r->tex_has_alpha[i] = texels_have_alpha(t.texels, (size_t)t.width * t.height) ? 1 : 0;
    }
    CUDA_TRY(r->texviews.upload(tv.data(), tv.size(), s));
    CUDA_TRY(cudaStreamSynchronize(s));
    r->h_texviews = tv;

    SceneView& sv = r->sv;
    sv.objects = r->objects.p; sv.matrices = r->matrices.p; sv.materials = r->materials.p; sv.lights = r->lights.p;
    sv.triangles = r->triangles.p; sv.positions = r->positions.p; sv.normals = r->normals.p;
    sv.textures = r->texviews.p; sv.light_count = (int32_t)scene->light_count; sv.material_count = (int32_t)scene->material_count; sv.cfg = scene->config;
    sv.scene_radius = scene_radius_of(scene->config);
    sv.alpha_blending = scene->enable_alpha_blending ? 1 : 0;
    update_shadow_retrace(r, scene->materials, scene->material_count);
    r->bv.t_min = scene->config.bvh_hit_min > 0 ? scene->config.bvh_hit_min : kEps;

    int rc = upload_bvh(r, scene, cache_path, cache_hit);
    if (rc != ATN_OK) return rc;

    r->camera = *camera;
    r->width = width; r->height = height;
    const size_t n = (size_t)width * height;
    // path buffers for a wavefront of about 32 M paths (four samples of a 4K frame, all sixteen of a 1080p one): up to
    // kMaxBatch samples are traced together (slot_info); ~250 B per path, 8 GB at the cap
    r->batch_cap = (int)std::max<size_t>(1, std::min<size_t>((size_t)ATN_MAX_BATCH, (r->wavefront_paths + n / 2) / n));
    r->path_capacity = n * (size_t)r->batch_cap;
    rc = alloc_paths(r, r->path_capacity, n);
    if (rc != ATN_OK) return rc;
    CUDA_TRY(r->image.resize(n));
    CUDA_TRY(cudaMemsetAsync(r->image.p, 0, n * sizeof(float4), s));
    if (r->seed_count == 0) {
        // default: every pixel seed 0 until atn_set_seeds is called
        std::vector<uint32_t> zero(1, 0u);
        CUDA_TRY(r->seeds.upload(zero.data(), 1, s));
        r->seed_count = 1;
    }
    CUDA_TRY(cudaStreamSynchronize(s));
    r->frame = 0;
    r->has_scene = true;
    return ATN_OK;
}

ATN_API int atn_upload_scene(AtnRenderer* r, const AtnSceneBlob* scene, const AtnCameraParameter* camera,
                             int32_t width, int32_t height)
{
    return upload_scene_impl(r, scene, camera, width, height, nullptr, nullptr);
}

ATN_API int atn_upload_scene_cached(AtnRenderer* r, const AtnSceneBlob* scene, const AtnCameraParameter* camera,
                                    int32_t width, int32_t height, const char* cache_path, int* cache_hit)
{
    if (!cache_path || !cache_path[0]) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_upload_scene_cached: no cache path");
    return upload_scene_impl(r, scene, camera, width, height, cache_path, cache_hit);
}

// ---- .sbvh files (sbvh::importTree / exportTree, sbvh.cpp:1220-1436) ------------------------------------------------
ATN_API int atn_sbvh_read(const char* path, int32_t offset_tri_idx, const char* const* material_names, uint32_t material_count,
                          AtnBvhNode** nodes, uint32_t* node_count, AtnSbvhInfo* info)
{
    if (!path || !nodes || !node_count) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_sbvh_read: NULL argument");
    SbvhFile f = read_sbvh(path, offset_tri_idx, material_names, material_count);
    if (!f.error.empty()) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_sbvh_read: %s", f.error.c_str());
    AtnBvhNode* out = static_cast<AtnBvhNode*>(malloc(std::max<size_t>(1, f.nodes.size()) * sizeof(AtnBvhNode)));
    if (!out) return set_error(ATN_ERR_TOO_LARGE, "atn_sbvh_read: out of host memory");
    if (!f.nodes.empty()) memcpy(out, f.nodes.data(), f.nodes.size() * sizeof(AtnBvhNode));
    *nodes = out; *node_count = (uint32_t)f.nodes.size();
    if (info) {
        memcpy(info->version, f.header.version, 4);
        info->max_depth = f.header.maxDepth; info->voxel_material_count = f.header.cntMtrlForVoxel;
        for (int a = 0; a < 3; a++) { info->boxmin[a] = f.header.boxmin[a]; info->boxmax[a] = f.header.boxmax[a]; }
    }
    return ATN_OK;
}

ATN_API int atn_sbvh_write(const char* path, const AtnBvhNode* nodes, uint32_t node_count, uint32_t max_depth,
                           const float boxmin[3], const float boxmax[3],
                           const int32_t* voxel_material_ids, const char* const* voxel_material_names, uint32_t voxel_material_count)
{
    if (!path || (node_count && !nodes) || !boxmin || !boxmax || (voxel_material_count && (!voxel_material_ids || !voxel_material_names)))
        return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_sbvh_write: NULL argument");
    const std::string err = write_sbvh(path, nodes, node_count, max_depth, boxmin, boxmax, voxel_material_ids, voxel_material_names, voxel_material_count);
    if (!err.empty()) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_sbvh_write: %s", err.c_str());
    return ATN_OK;
}

ATN_API void atn_free(void* p) { free(p); }

ATN_API int atn_update_camera(AtnRenderer* r, const AtnCameraParameter* camera)
{
    if (!r || !camera) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_update_camera: NULL argument");
    r->camera = *camera;
    return ATN_OK;
}

ATN_API int atn_update_top_layer(AtnRenderer* r, const AtnBvhNode* top_nodes, uint32_t node_count,
                                 const AtnObjectParameter* objects, uint32_t object_count,
                                 const AtnMat4* matrices, uint32_t matrix_count)
{
    if (!r || !top_nodes || !objects) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_update_top_layer: NULL argument");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_update_top_layer: no scene uploaded");
    CUDA_TRY(cudaSetDevice(r->device));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    // Only list 0 is encoded again, against the roots / stack needs / root boxes kept from the upload: the cost is that of
    // the top layer (instances), whatever the size of the bottom trees, like the reference's updateBVH (renderer.cpp:133-153).
    AtnSceneBlob b{};
    b.objects = objects; b.object_count = object_count;
    b.matrices = matrices; b.matrix_count = matrix_count;
    WideBvh top = encode_top(top_nodes, node_count, &b, r->bottom);
    if (!top.error.empty()) return set_error(ATN_ERR_INVALID_ARGUMENT, "top layer: %s", top.error.c_str());
    if (top.max_stack > kSmemStack + kLocalStackLarge)       // refused before anything on the device changes
        return set_error(ATN_ERR_TOO_LARGE, "BVH needs a traversal stack of %d entries (limit %d)", top.max_stack, kSmemStack + kLocalStackLarge);
    const size_t bottom_nodes = (size_t)r->bottom.node_count, total = bottom_nodes + top.nodes.size();
    if (total > r->node_capacity) {
        // more top-layer nodes than the array has room for: a larger array, bottom trees copied on the device
        DevBuf<QuadNode> grown;
        CUDA_TRY(grown.resize(total + top.nodes.size() + 64));
        CUDA_TRY(cudaMemcpyAsync(grown.p, r->nodes.p, bottom_nodes * sizeof(QuadNode), cudaMemcpyDeviceToDevice, r->stream));
        CUDA_TRY(cudaStreamSynchronize(r->stream));
        std::swap(grown.p, r->nodes.p); std::swap(grown.n, r->nodes.n);
        r->node_capacity = r->nodes.n;
    }
    CUDA_TRY(r->objects.upload(objects, object_count, r->stream));
    CUDA_TRY(r->matrices.upload(matrices, matrix_count, r->stream));
    r->sv.objects = r->objects.p; r->sv.matrices = r->matrices.p;
    {
        const int rc = upload_ref_top(r, top_nodes, node_count);
        if (rc != ATN_OK) return rc;
    }
    if (!top.nodes.empty())
        CUDA_TRY(cudaMemcpyAsync(r->nodes.p + bottom_nodes, top.nodes.data(), top.nodes.size() * sizeof(QuadNode), cudaMemcpyHostToDevice, r->stream));
    return bind_top_layer(r, total, top.insts, top.top_root, top.max_stack);
}

ATN_API int atn_update_materials(AtnRenderer* r, const AtnMaterialParameter* m, uint32_t count)
{
    if (!r || !m) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_update_materials: NULL argument");
    CUDA_TRY(cudaSetDevice(r->device));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    CUDA_TRY(r->materials.upload(m, count, r->stream));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    r->sv.materials = r->materials.p; r->sv.material_count = (int32_t)count;
    update_shadow_retrace(r, m, count);
    return ATN_OK;
}

ATN_API int atn_update_texture(AtnRenderer* r, uint32_t index, const AtnTextureDesc* texture)
{
    if (!r || !texture || !texture->texels) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_update_texture: NULL argument");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_update_texture: no scene uploaded");
    if (index >= r->h_texviews.size()) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_update_texture: index %u of %zu", index, r->h_texviews.size());
    const TexView& tv = r->h_texviews[index];
    if (!tv.texels || texture->width != tv.w || texture->height != tv.h)
        return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_update_texture: texture %u is %dx%d, got %dx%d (the size is fixed at upload)",
                         index, tv.w, tv.h, texture->width, texture->height);
    CUDA_TRY(cudaSetDevice(r->device));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    CUDA_TRY(cudaMemcpyAsync(const_cast<AtnVec4*>(tv.texels), texture->texels, (size_t)tv.w * tv.h * sizeof(AtnVec4),
                             cudaMemcpyHostToDevice, r->stream));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    if (index < r->tex_has_alpha.size()) {
        r->tex_has_alpha[index] = texels_have_alpha(texture->texels, (size_t)tv.w * tv.h) ? 1 : 0;
        if (r->tex_has_alpha[index]) r->alpha_surfaces = true;      // conservative until the next material update
    }
    return ATN_OK;
}

ATN_API int atn_update_lights(AtnRenderer* r, const AtnLightParameter* l, uint32_t count)
{
    if (!r || (!l && count)) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_update_lights: NULL argument");
    CUDA_TRY(cudaSetDevice(r->device));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    CUDA_TRY(r->lights.upload(l, count, r->stream));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    r->sv.lights = r->lights.p; r->sv.light_count = (int32_t)count;
    remember_lights(r, l, count);
    return refresh_light_bounds(r);
}

ATN_API int atn_update_config(AtnRenderer* r, const AtnSceneConfig* cfg)
{
    if (!r || !cfg) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_update_config: NULL argument");
    r->sv.cfg = *cfg;
    r->sv.scene_radius = scene_radius_of(*cfg);
    r->bv.t_min = cfg->bvh_hit_min > 0 ? cfg->bvh_hit_min : kEps;
    r->rv.t_min = r->bv.t_min;
    return ATN_OK;
}

ATN_API int atn_set_seeds(AtnRenderer* r, const uint32_t* seeds, uint32_t count)
{
    if (!r || !seeds || count == 0) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_set_seeds: NULL/empty");
    CUDA_TRY(cudaSetDevice(r->device));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    CUDA_TRY(r->seeds.upload(seeds, count, r->stream));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    r->seed_count = count;
    return ATN_OK;
}

ATN_API int atn_set_frame(AtnRenderer* r, uint32_t frame) { if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL"); r->frame = frame; return ATN_OK; }
ATN_API int atn_get_frame(AtnRenderer* r, uint32_t* frame) { if (!r || !frame) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL"); *frame = r->frame; return ATN_OK; }
ATN_API int atn_set_progressive(AtnRenderer* r, int enable) { if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL"); r->gather_mode = enable ? ATN_GATHER_PROGRESSIVE : ATN_GATHER_OVERWRITE; return ATN_OK; }
ATN_API int atn_set_gather_mode(AtnRenderer* r, int mode)
{
    if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL");
    if (mode != ATN_GATHER_OVERWRITE && mode != ATN_GATHER_PROGRESSIVE && mode != ATN_GATHER_SUM)
        return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_set_gather_mode: mode %d is none of ATN_GATHER_OVERWRITE / PROGRESSIVE / SUM", mode);
    r->gather_mode = mode;
    return ATN_OK;
}
ATN_API int atn_set_texture_fetch(AtnRenderer* r, int device_semantics)
{
    if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL");
    r->sv.device_texture_fetch = device_semantics ? 1 : 0;
    return ATN_OK;
}
ATN_API int atn_set_alpha_blending(AtnRenderer* r, int enable)
{
    if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL");
    r->sv.alpha_blending = enable ? 1 : 0;
    update_shadow_retrace(r, nullptr, 0);
    return ATN_OK;
}
ATN_API int atn_set_wavefront_paths(AtnRenderer* r, uint64_t max_paths)
{
    if (!r || max_paths == 0) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_set_wavefront_paths: NULL handle or zero paths");
    r->wavefront_paths = (size_t)std::min<uint64_t>(max_paths, (uint64_t)1 << 30);
    return ATN_OK;
}

ATN_API int atn_set_exact_ties(AtnRenderer* r, int enable) { if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL"); r->exact_ties = enable ? 1 : 0; return ATN_OK; }
ATN_API int atn_set_hit_distance_limit(AtnRenderer* r, float d) { if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL"); r->hit_distance_limit = fminf(d, kInf); return ATN_OK; }
ATN_API int atn_set_profiling(AtnRenderer* r, int enable) { if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL"); r->profiling = enable & 3; return ATN_OK; }
ATN_API int atn_get_stream(AtnRenderer* r, void** stream) { if (!r || !stream) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL"); *stream = (void*)r->stream; return ATN_OK; }

ATN_API int atn_reset(AtnRenderer* r)
{
    if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL");
    r->frame = 0;
    if (r->image.p) {
        CUDA_TRY(cudaSetDevice(r->device));
        CUDA_TRY(cudaMemsetAsync(r->image.p, 0, r->image.n * sizeof(float4), r->stream));
    }
    return ATN_OK;
}

ATN_API int atn_set_bounce_tap(AtnRenderer* r, AtnBounceRecord* records, int32_t bounce_capacity)
{
    if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL");
    r->tap_host = records;
    r->tap_cap = records ? bounce_capacity : 0;
    return ATN_OK;
}

ATN_API int atn_synchronize(AtnRenderer* r)
{
    if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "NULL");
    CUDA_TRY(cudaSetDevice(r->device));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    return ATN_OK;
}

// The bounce loop of one camera sample: closest hit, the three shade stages, shadow rays.
// `svgf_w2c` != nullptr: the SVGF shade variant (pb.aov_nd / pb.aov_am then point at the SVGF AOV set).
//
// Where the exact walk goes.  k_exact is a tail of a few thousand rays with hundreds of dependent steps each: 0.1 - 0.4 ms in
// which it keeps a tenth of the SMs' warp slots busy.  Run between the main kernels it costs that time twice per bounce (9 %
// of a frame); so it runs NEXT to them, on a second stream:
//
//   main:  trace closest b | surface  nee  bsdf (paths with a decided hit) | trace shadow b | trace closest b+1 | surface b+1 ...
//   side:                  | k_side b: exact walk + the three stages, its paths  |              | exact shadow b    |
//
// A path whose closest hit is with the exact walk is marked (kHitPending), left alone by the main shade kernels and shaded
// by the side pass (k_side: one launch, one thread per path), which appends to the same next-bounce and shadow queues; the
// main stream waits for the side pass before it traces the shadow rays.  The shadow rays of the exact walk only add to
// their own paths' contributions, into a buffer of their own (contrib_late) that the gather folds in: nothing waits for
// them before the gather.
// With the bounce tap on (parity runs) everything stays on one stream.
static cudaEvent_t side_event(AtnRenderer* r, size_t i)
{
    while (r->side_events.size() <= i) {
        cudaEvent_t ev = nullptr;
        if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        r->side_events.push_back(ev);
    }
    return r->side_events[i];
}

static int launch_bounces(AtnRenderer* r, const FrameParams& fp, const PathBuffers& pb, bool tap_now, const AtnMat4* svgf_w2c,
                          size_t* ev_i, uint64_t* launches, int first_bounce = 0)
{
    cudaStream_t s = r->stream, side = r->side_stream;
    AtnMat4 w2c;
    for (int i = 0; i < 16; i++) w2c.m[i] = svgf_w2c ? svgf_w2c->m[i] : 0.0f;
    const bool overlap = r->exact_ties && !tap_now && !r->shadow_retrace && !(r->exp_bits & 1);
    const int side_grid = r->sm_count * 4;
    size_t se = 0;
    cudaEvent_t shadow_done = nullptr;            // side stream: the exact walk of the previous bounce's shadow rays
    for (int b = first_bounce; b < fp.max_bounce; b++) {
        Counters* c = r->counters.p;
        const int32_t* q = (b == 0 && fp.tile == 0) ? nullptr : pb.queue[b & 1];
        const float tmax = (b == 0) ? kInf : r->hit_distance_limit;
        const ShadeQueues mainq{ q, &c->n_active[b], pb.q_lit, &c->n_lit[b], overlap ? b + 1 : 0, c->fetch_shade[b] };
        if (r->profiling & 1) CUDA_TRY(cudaEventRecord(r->trace_events[(*ev_i)++], s));
        ClosestPolicy cpol{ pb, q, tmax, b == 0 ? pb.hit_mesh : nullptr, (uint8_t)(overlap ? b + 1 : 0) };
        launch_trace<ClosestPolicy>(r, s, cpol, &c->n_active[b], 0, &c->fetch_closest[b], r->visits.p, pb.q_exact, &c->n_exact_closest[b], !overlap);
        if (r->profiling & 1) CUDA_TRY(cudaEventRecord(r->trace_events[(*ev_i)++], s));
        cudaEvent_t side_done = nullptr;
        if (overlap) {
            cudaEvent_t traced = side_event(r, se++); side_done = side_event(r, se++);
            if (!traced || !side_done) return set_error(ATN_ERR_CUDA, "cudaEventCreate failed");
            CUDA_TRY(cudaEventRecord(traced, s));
            CUDA_TRY(cudaStreamWaitEvent(side, traced, 0));
            if (svgf_w2c) k_side<true><<<side_grid, 128, 0, side>>>(r->rv, cpol, fp, r->camera, r->sv, r->seeds.p, pb, c, b, w2c);
            else k_side<false><<<side_grid, 128, 0, side>>>(r->rv, cpol, fp, r->camera, r->sv, r->seeds.p, pb, c, b, w2c);
            CUDA_TRY(cudaEventRecord(side_done, side));
            *launches += 1;
        }
        if (svgf_w2c) k_surface<false, true><<<r->surface_grid, kShadeBlock, 0, s>>>(fp, r->camera, r->sv, pb, mainq, b, nullptr, 0, w2c);
        else if (tap_now) k_surface<true, false><<<r->surface_grid, kShadeBlock, 0, s>>>(fp, r->camera, r->sv, pb, mainq, b, r->tap_dev.p, r->tap_cap, w2c);
        else k_surface<false, false><<<r->surface_grid, kShadeBlock, 0, s>>>(fp, r->camera, r->sv, pb, mainq, b, nullptr, 0, w2c);
        k_nee<<<r->nee_grid, kShadeBlock, 0, s>>>(fp, r->sv, r->seeds.p, pb, mainq, b);
        if (tap_now) k_bsdf<true><<<r->bsdf_grid, kShadeBlock, 0, s>>>(fp, r->sv, r->seeds.p, pb, c, mainq, b, r->tap_dev.p, r->tap_cap);
        else k_bsdf<false><<<r->bsdf_grid, kShadeBlock, 0, s>>>(fp, r->sv, r->seeds.p, pb, c, mainq, b, nullptr, 0);
        if (side_done) CUDA_TRY(cudaStreamWaitEvent(s, side_done, 0));       // the shadow and next-bounce queues are complete
        if (r->profiling & 1) CUDA_TRY(cudaEventRecord(r->trace_events[(*ev_i)++], s));
        if (r->shadow_retrace) {
            // HitTestToTargetLight's loop: lookup 0 over the shadow queue, lookups 1 .. 9 over the rays that restarted behind
            // an ignored hit (a handful; an empty pass is a few microseconds).  One stream, exact walk in line.
            for (int pass = 0; pass < 10; pass++) {
                const int32_t* q_in = pass == 0 ? pb.q_shadow : pb.q_retrace[(pass - 1) & 1];
                const int32_t* n_in = pass == 0 ? &c->n_shadow[b] : &c->retrace[b].n[pass - 1];
                int32_t* fetch = pass == 0 ? &c->fetch_shadow[b] : &c->retrace[b].fetch[pass - 1];
                int32_t* n_ex = pass == 0 ? &c->n_exact_shadow[b] : &c->retrace[b].n_exact[pass - 1];
                if (tap_now) {
                    ShadowPolicy<true, true> pol{ pb, r->sv, r->tap_dev.p, r->tap_cap, b, q_in, pb.q_retrace[pass & 1], &c->retrace[b].n[pass] };
                    launch_trace<ShadowPolicy<true, true>>(r, s, pol, n_in, 0, fetch, r->visits.p, pb.q_exact_s, n_ex);
                }
                else {
                    ShadowPolicy<false, true> pol{ pb, r->sv, nullptr, 0, b, q_in, pb.q_retrace[pass & 1], &c->retrace[b].n[pass] };
                    launch_trace<ShadowPolicy<false, true>>(r, s, pol, n_in, 0, fetch, r->visits.p, pb.q_exact_s, n_ex);
                }
            }
            *launches += 9 * (r->exact_ties ? 2 : 1);
        }
        else if (tap_now) {
            ShadowPolicy<true> pol{ pb, r->sv, r->tap_dev.p, r->tap_cap, b, pb.q_shadow, nullptr, nullptr };
            launch_trace<ShadowPolicy<true>>(r, s, pol, &c->n_shadow[b], 0, &c->fetch_shadow[b], r->visits.p, pb.q_exact_s, &c->n_exact_shadow[b]);
        }
        else {
            ShadowPolicy<false> pol{ pb, r->sv, nullptr, 0, b, pb.q_shadow, nullptr, nullptr };
            launch_trace<ShadowPolicy<false>>(r, s, pol, &c->n_shadow[b], 0, &c->fetch_shadow[b], r->visits.p, pb.q_exact_s, &c->n_exact_shadow[b], !overlap);
            if (overlap) {
                cudaEvent_t traced = side_event(r, se++); shadow_done = side_event(r, se++);
                if (!traced || !shadow_done) return set_error(ATN_ERR_CUDA, "cudaEventCreate failed");
                CUDA_TRY(cudaEventRecord(traced, s));
                CUDA_TRY(cudaStreamWaitEvent(side, traced, 0));
                launch_exact<ShadowPolicy<false>>(r, side, pol, pb.q_exact_s, &c->n_exact_shadow[b]);
                CUDA_TRY(cudaEventRecord(shadow_done, side));
                *launches += 1;
            }
        }
        if (r->profiling & 1) CUDA_TRY(cudaEventRecord(r->trace_events[(*ev_i)++], s));
        *launches += 5 + (r->exact_ties && !overlap ? 2 : 0);
    }
    if (shadow_done) CUDA_TRY(cudaStreamWaitEvent(s, shadow_done, 0));       // the gather reads every contribution
    return ATN_OK;
}

// atn_render_region and atn_render_tiles: `tile` = 0 renders the rectangle, `tile` > 0 the interleaved tile set of
// (tile_world, tile_rank) as ONE wavefront (x0 .. y1 ignored).
static int render_impl(AtnRenderer* r, int32_t width, int32_t height, int32_t x0, int32_t y0, int32_t x1, int32_t y1,
                       int32_t tile, int32_t tile_world, int32_t tile_rank,
                       int32_t sample_begin, int32_t sample_end, int32_t max_bounce, int32_t rr_bounce)
{
    if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render: NULL handle");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_render: no scene uploaded");
    if (width != r->width || height != r->height)
        return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render: size %dx%d differs from the uploaded %dx%d", width, height, r->width, r->height);
    if (max_bounce < 1 || max_bounce > kMaxBounce) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render: max_bounce %d not in [1,%d]", max_bounce, kMaxBounce);
    if (x0 < 0 || y0 < 0 || x1 > width || y1 > height || x0 >= x1 || y0 >= y1) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render: bad region");
    if (sample_end < sample_begin) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render: bad sample range");
    if (rr_bounce > max_bounce) rr_bounce = max_bounce - 1;       // pathtracing.cpp (host) :281-283
    CUDA_TRY(cudaSetDevice(r->device));
    cudaStream_t s = r->stream;

    FrameParams fp;
    fp.width = width; fp.height = height; fp.x0 = x0; fp.y0 = y0; fp.rw = x1 - x0; fp.rh = y1 - y0;
    fp.max_bounce = max_bounce; fp.rr_bounce = rr_bounce; fp.seed_count = r->seed_count;
    fp.tile = 0; fp.tiles_x = 0; fp.tile_world = 1; fp.tile_rank = 0; fp.swizzle = 0;
    if (tile > 0) {
        const int tiles_x = (width + tile - 1) / tile, tiles_y = (height + tile - 1) / tile;
        const int total = tiles_x * tiles_y;
        const int mine = total > tile_rank ? (total - tile_rank + tile_world - 1) / tile_world : 0;
        fp.tile = tile; fp.tiles_x = tiles_x; fp.tile_world = tile_world; fp.tile_rank = tile_rank;
        fp.x0 = 0; fp.y0 = 0; fp.rw = tile * tile; fp.rh = mine;
        if (mine == 0) return ATN_OK;                                  // more ranks than tiles: nothing to do
        if (r->tap_host) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_tiles: the bounce tap is not available on a tile set");
        if ((size_t)fp.rw * fp.rh > r->path_capacity) return set_error(ATN_ERR_TOO_LARGE, "atn_render_tiles: tile set larger than the path buffers");
    }
    const int n = fp.rw * fp.rh;
    uint64_t n_pixels = (uint64_t)n;                 // pixels inside the frame (an edge tile is partly outside)
    if (tile > 0) {
        n_pixels = 0;
        for (int j = 0; j < fp.rh; j++) {
            const int g = j * tile_world + tile_rank, tx = g % fp.tiles_x, ty = g / fp.tiles_x;
            n_pixels += (uint64_t)std::min(tile, width - tx * tile) * (uint64_t)std::min(tile, height - ty * tile);
        }
    }
    const bool tap = r->tap_host != nullptr;
#if ATN_PIXEL_BLOCKS
    fp.swizzle = (!tap && tile == 0 && fp.rw % 8 == 0 && fp.rh % 4 == 0) ? 1 : 0;      // the tap's records are indexed by pixel
#endif
    if (tap) {
        CUDA_TRY(r->tap_dev.resize((size_t)n * r->tap_cap));
        if (r->tap_cap < max_bounce) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render: tap capacity %d < max_bounce %d", r->tap_cap, max_bounce);
    }
    const int spp = sample_end - sample_begin;
    const size_t n_ev = (r->profiling & 1) ? (size_t)spp * max_bounce * 4 : 0;
    while (r->trace_events.size() < n_ev) { cudaEvent_t ev; CUDA_TRY(cudaEventCreate(&ev)); r->trace_events.push_back(ev); }
    size_t ev_i = 0;
    uint64_t launches = 0;
    const int gen_grid = std::min((n + 255) / 256, r->sm_count * 8);

    CUDA_TRY(cudaMemsetAsync(r->totals.p, 0, sizeof(Totals), s));
    CUDA_TRY(cudaMemsetAsync(r->visits.p, 0, sizeof(Visits), s));
    CUDA_TRY(cudaEventRecord(r->ev0, s));
    // Samples are traced `batch` at a time as one wavefront (slot_info): as many as the path buffers hold, one when
    // the bounce tap is on (its records are indexed by pixel).
    const int batch_max = tap ? 1 : std::max<int>(1, (int)std::min<size_t>((size_t)r->batch_cap, r->path_capacity / (size_t)n));
    for (int smp = sample_begin; smp < sample_end;) {
        const int batch = std::min(batch_max, sample_end - smp);
        fp.frame = r->frame + (uint32_t)smp;
        fp.batch = batch;
        const int paths = n * batch;
        const int grid_b = std::min((paths + 255) / 256, r->sm_count * 8);
        CUDA_TRY(cudaMemsetAsync(r->counters.p, 0, sizeof(Counters), s));
        const bool tap_now = tap && smp == sample_end - 1;
        if (tap_now) CUDA_TRY(cudaMemsetAsync(r->tap_dev.p, 0, (size_t)n * r->tap_cap * sizeof(AtnBounceRecord), s));
        k_generate<<<grid_b, 256, 0, s>>>(fp, r->camera, r->seeds.p, r->pb, r->counters.p);
        launches++;
        const int rc = launch_bounces(r, fp, r->pb, tap_now, nullptr, &ev_i, &launches);
        if (rc != ATN_OK) return rc;
        k_gather<<<gen_grid, 256, 0, s>>>(fp, r->pb, r->image.p, r->gather_mode, smp == sample_begin ? 1 : 0,
                                          smp + batch >= sample_end ? 1 : 0, r->counters.p, r->totals.p);
        launches++;
        smp += batch;
    }
    // Exact ray counts: k_gather folds every sample's queue counters into Totals on the device; they are read on demand
    // (atn_get_stats).  The call returns without waiting for the GPU unless the bounce tap has to be copied out.
    const int rc_end = end_call(r, AtnRenderer::PEND_RENDER, launches, n_pixels * (uint64_t)spp, ev_i);
    if (rc_end != ATN_OK) return rc_end;
    if (tap) {
        CUDA_TRY(cudaMemcpyAsync(r->tap_host, r->tap_dev.p, (size_t)n * r->tap_cap * sizeof(AtnBounceRecord), cudaMemcpyDeviceToHost, s));
        CUDA_TRY(cudaStreamSynchronize(s));
    }
    return ATN_OK;
}

ATN_API int atn_render_region(AtnRenderer* r, int32_t width, int32_t height, int32_t x0, int32_t y0, int32_t x1, int32_t y1,
                              int32_t sample_begin, int32_t sample_end, int32_t max_bounce, int32_t rr_bounce)
{
    return render_impl(r, width, height, x0, y0, x1, y1, 0, 1, 0, sample_begin, sample_end, max_bounce, rr_bounce);
}

ATN_API int atn_render_tiles(AtnRenderer* r, int32_t width, int32_t height, int32_t tile, int32_t world, int32_t rank,
                             int32_t sample_begin, int32_t sample_end, int32_t max_bounce, int32_t rr_bounce)
{
    if (tile < 8 || tile > 1024 || world < 1 || rank < 0 || rank >= world)
        return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_tiles: tile %d (8..1024), world %d, rank %d", tile, world, rank);
    return render_impl(r, width, height, 0, 0, width, height, tile, world, rank, sample_begin, sample_end, max_bounce, rr_bounce);
}

ATN_API int atn_render_ao(AtnRenderer* r, int32_t width, int32_t height, int32_t spp, int32_t num_rays, float radius)
{
    if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_ao: NULL handle");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_render_ao: no scene uploaded");
    if (width != r->width || height != r->height)
        return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_ao: size %dx%d differs from the uploaded %dx%d", width, height, r->width, r->height);
    if (spp < 1 || num_rays < 1 || !(radius > 0.0f))
        return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_ao: spp %d, num_rays %d, radius %g must be positive", spp, num_rays, radius);
    CUDA_TRY(cudaSetDevice(r->device));
    cudaStream_t s = r->stream;
    FrameParams fp;
    fp.width = width; fp.height = height; fp.x0 = 0; fp.y0 = 0; fp.rw = width; fp.rh = height; fp.tile = 0; fp.tiles_x = 0; fp.tile_world = 1; fp.tile_rank = 0; fp.swizzle = 0;
    fp.max_bounce = 1; fp.rr_bounce = 1; fp.seed_count = r->seed_count; fp.batch = 1;
    const int n = width * height;
    const int grid = std::min((n + 255) / 256, r->sm_count * 8);
    const int sgrid = r->sm_count * 8;
    uint64_t launches = 0;
    CUDA_TRY(cudaMemsetAsync(r->totals.p, 0, sizeof(Totals), s));
    CUDA_TRY(cudaMemsetAsync(r->visits.p, 0, sizeof(Visits), s));
    CUDA_TRY(cudaEventRecord(r->ev0, s));
    // samples share a wavefront as in the path tracer (slot_info): the launches of a 2 M-ray sample are short enough for
    // their tails and the exact walk's latency to show
    const int batch_max = std::max<int>(1, (int)std::min<size_t>((size_t)r->batch_cap, r->path_capacity / (size_t)n));
    for (int smp = 0; smp < spp;) {
        const int batch = std::min(batch_max, spp - smp);
        fp.frame = r->frame + (uint32_t)smp;
        fp.batch = batch;
        const int grid_b = std::min((n * batch + 255) / 256, r->sm_count * 8);
        Counters* c = r->counters.p;
        CUDA_TRY(cudaMemsetAsync(c, 0, sizeof(Counters), s));
        k_generate<<<grid_b, 256, 0, s>>>(fp, r->camera, r->seeds.p, r->pb, c);
        {
            ClosestPolicy pol{ r->pb, nullptr, kInf, nullptr, 0 };
            launch_trace<ClosestPolicy>(r, s, pol, &c->n_active[0], 0, &c->fetch_closest[0], r->visits.p, r->pb.q_exact, &c->n_exact_closest[0]);
        }
        k_ao_begin<<<sgrid, kShadeBlock, 0, s>>>(fp, r->sv, r->seeds.p, r->pb, c);
        launches += 3;
        for (int i = 0; i < num_rays; i++) {
            k_ao_ray<<<sgrid, kShadeBlock, 0, s>>>(fp, r->seeds.p, r->pb, c);
            CUDA_TRY(cudaMemsetAsync(&c->fetch_query, 0, sizeof(int32_t), s));
            CUDA_TRY(cudaMemsetAsync(&c->n_exact_query, 0, sizeof(int32_t), s));
            ClosestPolicy pol{ r->pb, r->pb.q_lit, radius, nullptr, 0 };
            launch_trace<ClosestPolicy>(r, s, pol, &c->n_lit[0], 0, &c->fetch_query, r->visits.p, r->pb.q_exact, &c->n_exact_query);
            if (r->alpha_surfaces) CUDA_TRY(cudaMemsetAsync(&c->retrace[0], 0, sizeof(Counters::Retrace), s));
            k_ao_accum<<<sgrid, kShadeBlock, 0, s>>>(r->sv, r->pb, r->pb.q_lit, &c->n_lit[0], r->pb.q_retrace[0], &c->retrace[0].n[0], radius,
                                                     r->alpha_surfaces ? 0 : 1);
            launches += 3;
            // the AO ray's loop through surfaces that are translucent by alpha: lookups 2 .. 10 over the rays that restarted
            for (int pass = 1; r->alpha_surfaces && pass < 10; pass++) {
                const int32_t* q_in = r->pb.q_retrace[(pass - 1) & 1];
                ClosestPolicy rp{ r->pb, q_in, radius, nullptr, 0 };
                launch_trace<ClosestPolicy>(r, s, rp, &c->retrace[0].n[pass - 1], 0, &c->retrace[0].fetch[pass - 1], r->visits.p, r->pb.q_exact,
                                            &c->retrace[0].n_exact[pass - 1]);
                k_ao_accum<<<sgrid, kShadeBlock, 0, s>>>(r->sv, r->pb, q_in, &c->retrace[0].n[pass - 1], r->pb.q_retrace[pass & 1],
                                                         &c->retrace[0].n[pass], radius, pass == 9 ? 1 : 0);
                launches += 2;
            }
        }
        k_ao_finish<<<sgrid, kShadeBlock, 0, s>>>(r->pb, c, num_rays);
        launches++;
        // n_active[0] primary rays + num_rays * n_lit[0] AO rays: folded into Totals by k_gather through
        // n_active[0] / n_shadow[0]; the AO rays are reported as shadow rays
        k_gather<<<grid, 256, 0, s>>>(fp, r->pb, r->image.p, r->gather_mode, smp == 0 ? 1 : 0, smp + batch >= spp ? 1 : 0, c, r->totals.p);
        launches++;
        smp += batch;
        r->pend_ao_batch = (uint64_t)batch;
    }
    r->pend_ao_factor = (uint64_t)num_rays * (uint64_t)spp;
    const int rc_end = end_call(r, AtnRenderer::PEND_AO, launches, (uint64_t)n * (uint64_t)spp, 0);
    if (rc_end != ATN_OK) return rc_end;
    r->frame += (uint32_t)spp;
    return ATN_OK;
}

// ---- N1: SVGF ----------------------------------------------------------------------------------
static SvgfAov svgf_aov_of(AtnRenderer* r, int k)
{
    AtnVec4* base = reinterpret_cast<AtnVec4*>(r->svgf_planes.p);
    const size_t n = r->svgf_pixels;
    SvgfAov a;
    a.normal_depth = base + (size_t)(k * 4 + 0) * n; a.albedo_meshid = base + (size_t)(k * 4 + 1) * n;
    a.color_variance = base + (size_t)r->svgf_cv_plane[k] * n; a.moment_tw = base + (size_t)(k * 4 + 3) * n;
    return a;
}

ATN_API int atn_set_motion_depth(AtnRenderer* r, const AtnVec4* motion_depth, size_t pixel_count)
{
    if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_set_motion_depth: NULL handle");
    if (!motion_depth) { r->svgf_has_motion = false; return ATN_OK; }
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_set_motion_depth: no scene uploaded");
    if (pixel_count != (size_t)r->width * r->height) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_set_motion_depth: pixel_count mismatch");
    CUDA_TRY(cudaSetDevice(r->device));
    CUDA_TRY(r->svgf_motion.upload(reinterpret_cast<const float4*>(motion_depth), pixel_count, r->stream));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    r->svgf_has_motion = true;
    return ATN_OK;
}

ATN_API int atn_render_svgf(AtnRenderer* r, int32_t width, int32_t height, int32_t max_bounce, int32_t rr_bounce, int32_t atrous_iters)
{
    if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_svgf: NULL handle");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_render_svgf: no scene uploaded");
    if (width != r->width || height != r->height)
        return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_svgf: size %dx%d differs from the uploaded %dx%d", width, height, r->width, r->height);
    if (max_bounce < 1 || max_bounce > kMaxBounce) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_svgf: max_bounce %d not in [1,%d]", max_bounce, kMaxBounce);
    if (atrous_iters <= 0) atrous_iters = 5;                      // SVGFParams::atrous_iter_cnt, svgf_types.h:74
    if (atrous_iters > 16) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_svgf: atrous_iters %d > 16", atrous_iters);
    if (rr_bounce > max_bounce) rr_bounce = max_bounce - 1;       // svgf.cpp:432-434
    if (r->tap_host) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_svgf: the bounce tap is not available on this renderer");
    CUDA_TRY(cudaSetDevice(r->device));
    cudaStream_t s = r->stream;
    const size_t n = (size_t)width * height;
    const int grid = (int)std::min<size_t>((n + 255) / 256, (size_t)r->sm_count * 8);

    const bool first_frame = r->frame == 0;
    if (r->svgf_pixels != n || !r->svgf_planes.p) {
        if (!first_frame) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_svgf: no history for frame %u (start at frame 0: atn_set_frame(0))", r->frame);
        CUDA_TRY(r->svgf_planes.resize(12 * n));
        r->svgf_pixels = n;
        if (r->svgf_tex) { cudaDestroyTextureObject(r->svgf_tex); r->svgf_tex = 0; }
        if (ATN_SVGF_TEX && 12 * n <= ((size_t)1 << 27)) {      // tex1Dfetch addresses 2^27 texels; larger frames use LDG only
            cudaResourceDesc rd{}; rd.resType = cudaResourceTypeLinear; rd.res.linear.devPtr = r->svgf_planes.p;
            rd.res.linear.desc = cudaCreateChannelDesc<float4>(); rd.res.linear.sizeInBytes = 12 * n * sizeof(float4);
            cudaTextureDesc td{}; td.readMode = cudaReadModeElementType;
            CUDA_TRY(cudaCreateTextureObject(&r->svgf_tex, &rd, &td, nullptr));
        }
    }
    if (first_frame) {
        r->svgf_cv_plane[0] = 2; r->svgf_cv_plane[1] = 6; r->svgf_scratch_plane = 11;
        k_fill_vec4<<<r->sm_count * 8, 256, 0, s>>>(r->svgf_planes.p, 12 * n, make_float4(0.0f, 0.0f, 0.0f, 1.0f));
    }
    if (r->svgf_has_motion && r->svgf_motion.n < n) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_svgf: motion buffer smaller than the image");

    FrameParams fp;
    fp.width = width; fp.height = height; fp.x0 = 0; fp.y0 = 0; fp.rw = width; fp.rh = height; fp.tile = 0; fp.tiles_x = 0; fp.tile_world = 1; fp.tile_rank = 0; fp.swizzle = 0;
    fp.max_bounce = max_bounce; fp.rr_bounce = rr_bounce; fp.seed_count = r->seed_count; fp.batch = 1;
    fp.frame = r->frame;
    const int cur_k = (int)(r->frame & 1u);
    AtnVec4* base = reinterpret_cast<AtnVec4*>(r->svgf_planes.p);
    SvgfView sv;
    sv.cur = svgf_aov_of(r, cur_k); sv.prev = svgf_aov_of(r, 1 - cur_k);
    sv.atrous[0] = base + 8 * n; sv.atrous[1] = base + 9 * n; sv.temporary = base + 10 * n;
    sv.scratch = base + (size_t)r->svgf_scratch_plane * n;
    sv.motion_depth = r->svgf_has_motion ? reinterpret_cast<const AtnVec4*>(r->svgf_motion.p) : nullptr;
    PathBuffers pb = r->pb;
    pb.aov_nd = reinterpret_cast<float4*>(sv.cur.normal_depth);
    pb.aov_am = reinterpret_cast<float4*>(sv.cur.albedo_meshid);
    const AtnMat4 w2c = camera_w2c(r->camera);
    const float camera_distance = screen_distance(r->camera, height);

    const size_t n_ev = (r->profiling & 1) ? (size_t)max_bounce * 4 : 0;
    while (r->trace_events.size() < n_ev) { cudaEvent_t ev; CUDA_TRY(cudaEventCreate(&ev)); r->trace_events.push_back(ev); }
    size_t ev_i = 0;
    uint64_t launches = 0;
    CUDA_TRY(cudaMemsetAsync(r->totals.p, 0, sizeof(Totals), s));
    CUDA_TRY(cudaMemsetAsync(r->visits.p, 0, sizeof(Visits), s));
    CUDA_TRY(cudaEventRecord(r->ev0, s));
    CUDA_TRY(cudaMemsetAsync(r->counters.p, 0, sizeof(Counters), s));
    k_generate<<<grid, 256, 0, s>>>(fp, r->camera, r->seeds.p, pb, r->counters.p);
    launches++;
    const int rc = launch_bounces(r, fp, pb, false, &w2c, &ev_i, &launches);
    if (rc != ATN_OK) return rc;
    k_svgf_prepare<<<grid, 256, 0, s>>>(fp, pb, sv, first_frame ? 1 : 0, r->counters.p, r->totals.p);
    if (!first_frame) { k_svgf_temporal<<<grid, 256, 0, s>>>(fp, pb, sv); launches++; }
    k_svgf_variance<<<grid, 256, 0, s>>>(fp, sv, camera_distance);
    // the variance pass wrote the filtered colour history to the scratch plane: that plane is the AOV set's from here on
    std::swap(r->svgf_cv_plane[cur_k], r->svgf_scratch_plane);
    sv.cur = svgf_aov_of(r, cur_k);
    sv.scratch = base + (size_t)r->svgf_scratch_plane * n;
    for (int i = 0; i < atrous_iters; i++)
        k_svgf_atrous<<<grid, 256, 0, s>>>(fp, sv, camera_distance, i, atrous_iters, r->image.p, r->svgf_tex, base);
    k_svgf_copy<<<grid, 256, 0, s>>>(fp, sv);
    launches += 3 + (uint64_t)atrous_iters;
    const int rc_end = end_call(r, AtnRenderer::PEND_RENDER, launches, (uint64_t)n, ev_i);
    if (rc_end != ATN_OK) return rc_end;
    r->frame += 1u;
    return ATN_OK;
}

ATN_API int atn_readback_svgf(AtnRenderer* r, AtnVec4* normal_depth, AtnVec4* albedo_meshid, AtnVec4* color_variance,
                              AtnVec4* moment_temporalweight, size_t pixel_count)
{
    if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_readback_svgf: NULL handle");
    if (!r->svgf_planes.p || r->frame == 0) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_readback_svgf: no SVGF frame rendered");
    if (pixel_count != r->svgf_pixels) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_readback_svgf: pixel_count mismatch");
    CUDA_TRY(cudaSetDevice(r->device));
    const SvgfAov a = svgf_aov_of(r, (int)((r->frame - 1u) & 1u));      // the set of the last frame rendered
    const AtnVec4* src[4] = { a.normal_depth, a.albedo_meshid, a.color_variance, a.moment_tw };
    AtnVec4* dst[4] = { normal_depth, albedo_meshid, color_variance, moment_temporalweight };
    for (int k = 0; k < 4; k++)
        if (dst[k]) CUDA_TRY(cudaMemcpyAsync(dst[k], src[k], pixel_count * sizeof(AtnVec4), cudaMemcpyDeviceToHost, r->stream));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    return ATN_OK;
}

// ---- N1: ReSTIR --------------------------------------------------------------------------------
__global__ void k_restir_init(Reservoir* res0, Reservoir* res1, RestirInfo* info0, RestirInfo* info1, size_t n)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        Reservoir r;
        r.w_sum = 0.0f; r.M = 0; r.y = -1; r.W = 0.0f; r.target_pdf = 0.0f;
        r.ls.pos = V3(0.0f); r.ls.dir = V3(0.0f); r.ls.dist = 0.0f; r.ls.nml = V3(0.0f); r.ls.color = V3(0.0f); r.ls.pdf = 0.0f; r.ls.attrib = 0;
        RestirInfo f;
        f.clear(); f.point_to_light = V3(0.0f); f.p = V3(0.0f); f.throughput = V3(0.0f);
        res0[i] = r; res1[i] = r; info0[i] = f; info1[i] = f;
    }
}

ATN_API int atn_render_restir(AtnRenderer* r, int32_t width, int32_t height, int32_t max_bounce, int32_t rr_bounce)
{
    if (!r) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_restir: NULL handle");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_render_restir: no scene uploaded");
    if (width != r->width || height != r->height)
        return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_restir: size %dx%d differs from the uploaded %dx%d", width, height, r->width, r->height);
    if (max_bounce < 1 || max_bounce > kMaxBounce) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_restir: max_bounce %d not in [1,%d]", max_bounce, kMaxBounce);
    if (rr_bounce > max_bounce) rr_bounce = max_bounce - 1;       // restir.cpp (host) :333-335
    if (r->tap_host) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_restir: the bounce tap is not available on this renderer");
    if (r->sv.light_count <= 0) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_restir: the scene has no light");
    CUDA_TRY(cudaSetDevice(r->device));
    cudaStream_t s = r->stream;
    const size_t n = (size_t)width * height;
    const int grid = (int)std::min<size_t>((n + 255) / 256, (size_t)r->sm_count * 8);
    const int sgrid = r->sm_count * 8;

    if (r->restir_pixels != n || !r->restir_res[0].p) {
        if (r->frame != 0) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_restir: no history for frame %u (start at frame 0: atn_set_frame(0))", r->frame);
        for (int k = 0; k < 2; k++) { CUDA_TRY(r->restir_res[k].resize(n)); CUDA_TRY(r->restir_info[k].resize(n)); }
        r->restir_pixels = n;
    }
    if (r->frame == 0) {
        k_restir_init<<<sgrid, 256, 0, s>>>(r->restir_res[0].p, r->restir_res[1].p, r->restir_info[0].p, r->restir_info[1].p, n);
        CUDA_TRY(cudaMemsetAsync(r->sh_c.p, 0, n * sizeof(float4), s));          // ShadowRay::lightcontrib of a new sequence
        r->restir_cur = 0;
    }
    if (r->svgf_has_motion && r->svgf_motion.n < n) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render_restir: motion buffer smaller than the image");

    FrameParams fp;
    fp.width = width; fp.height = height; fp.x0 = 0; fp.y0 = 0; fp.rw = width; fp.rh = height; fp.tile = 0; fp.tiles_x = 0; fp.tile_world = 1; fp.tile_rank = 0; fp.swizzle = 0;
    fp.max_bounce = max_bounce; fp.rr_bounce = rr_bounce; fp.seed_count = r->seed_count; fp.batch = 1;
    fp.frame = r->frame;
    const int cur = r->restir_cur, other = (cur + 1) & 1;
    RestirView rv;
    rv.cur = r->restir_res[cur].p; rv.other = r->restir_res[other].p;
    rv.info_cur = r->restir_info[cur].p; rv.info_other = r->restir_info[other].p;
    rv.motion_depth = r->svgf_has_motion ? reinterpret_cast<const AtnVec4*>(r->svgf_motion.p) : nullptr;
    const PathBuffers& pb = r->pb;
    const AtnMat4 w2c = camera_w2c(r->camera);
    Counters* c = r->counters.p;

    const size_t n_ev = (r->profiling & 1) ? (size_t)max_bounce * 4 : 0;
    while (r->trace_events.size() < n_ev) { cudaEvent_t ev; CUDA_TRY(cudaEventCreate(&ev)); r->trace_events.push_back(ev); }
    size_t ev_i = 0;
    uint64_t launches = 0;
    CUDA_TRY(cudaMemsetAsync(r->totals.p, 0, sizeof(Totals), s));
    CUDA_TRY(cudaMemsetAsync(r->visits.p, 0, sizeof(Visits), s));
    CUDA_TRY(cudaEventRecord(r->ev0, s));
    CUDA_TRY(cudaMemsetAsync(c, 0, sizeof(Counters), s));
    k_generate<<<grid, 256, 0, s>>>(fp, r->camera, r->seeds.p, pb, c);
    // bounce 0
    {
        ClosestPolicy pol{ pb, nullptr, kInf, pb.hit_mesh, 0 };
        launch_trace<ClosestPolicy>(r, s, pol, &c->n_active[0], 0, &c->fetch_closest[0], r->visits.p, pb.q_exact, &c->n_exact_closest[0]);
    }
    k_restir_begin<<<sgrid, kShadeBlock, 0, s>>>(fp, r->camera, r->sv, pb, rv, c, w2c);
    k_restir_candidate<<<sgrid, kShadeBlock, 0, s>>>(fp, r->sv, r->seeds.p, pb, rv, c);
    k_bsdf<false><<<r->bsdf_grid, kShadeBlock, 0, s>>>(fp, r->sv, r->seeds.p, pb, c, ShadeQueues{ nullptr, &c->n_active[0], pb.q_lit, &c->n_lit[0], 0, c->fetch_shade[0] }, 0, nullptr, 0);
    k_restir_visibility<<<sgrid, kShadeBlock, 0, s>>>(fp, pb, rv, c);
    {
        RestirShadowPolicy pol{ ShadowPolicy<false>{ pb, r->sv, nullptr, 0, 0, pb.q_shadow, nullptr, nullptr }, rv.cur };
        launch_trace<RestirShadowPolicy>(r, s, pol, &c->n_shadow[0], 0, &c->fetch_shadow[0], r->visits.p, pb.q_exact, &c->n_exact_shadow[0]);
    }
    launches += 7;
    if (r->frame > 1) { k_restir_temporal<<<sgrid, kShadeBlock, 0, s>>>(fp, r->sv, r->seeds.p, pb, rv); launches++; }
    k_restir_spatial<<<sgrid, kShadeBlock, 0, s>>>(fp, r->sv, r->seeds.p, pb, rv);
    k_restir_color<<<sgrid, kShadeBlock, 0, s>>>(fp, r->sv, pb, rv);
    launches += 2;
    r->restir_target = other;
    r->restir_cur = other;                                   // ReuseParams::Update
    // bounces > 0: the path tracer's
    const int rc = launch_bounces(r, fp, pb, false, nullptr, &ev_i, &launches, 1);
    if (rc != ATN_OK) return rc;
    // one frame of one sample per call: in overwrite mode the image holds this frame alone
    k_gather<<<grid, 256, 0, s>>>(fp, pb, r->image.p, r->gather_mode, 1, 1, c, r->totals.p);
    launches++;
    const int rc_end = end_call(r, AtnRenderer::PEND_RENDER, launches, (uint64_t)n, ev_i);
    if (rc_end != ATN_OK) return rc_end;
    r->frame += 1u;
    return ATN_OK;
}

ATN_API int atn_readback_restir(AtnRenderer* r, float* reservoirs, size_t pixel_count)
{
    if (!r || !reservoirs) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_readback_restir: NULL argument");
    if (!r->restir_res[0].p || r->frame == 0) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_readback_restir: no ReSTIR frame rendered");
    if (pixel_count != r->restir_pixels) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_readback_restir: pixel_count mismatch");
    CUDA_TRY(cudaSetDevice(r->device));
    std::vector<Reservoir> host(pixel_count);
    CUDA_TRY(cudaMemcpyAsync(host.data(), r->restir_res[r->restir_target].p, pixel_count * sizeof(Reservoir), cudaMemcpyDeviceToHost, r->stream));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    for (size_t i = 0; i < pixel_count; i++) {
        float* o = reservoirs + i * 5;
        o[0] = host[i].w_sum; o[1] = (float)host[i].M; o[2] = (float)host[i].y; o[3] = host[i].W; o[4] = host[i].target_pdf;
    }
    return ATN_OK;
}

ATN_API int atn_render(AtnRenderer* r, int32_t width, int32_t height, int32_t spp, int32_t max_bounce, int32_t rr_bounce)
{
    if (spp < 1) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_render: spp %d < 1", spp);
    const int rc = atn_render_region(r, width, height, 0, 0, width, height, 0, spp, max_bounce, rr_bounce);
    if (rc == ATN_OK) r->frame += (uint32_t)spp;
    return rc;
}

ATN_API int atn_readback(AtnRenderer* r, AtnVec4* host_image, size_t pixel_count)
{
    if (!r || !host_image) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_readback: NULL argument");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_readback: no scene uploaded");
    if (pixel_count != (size_t)r->width * r->height) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_readback: pixel_count mismatch");
    CUDA_TRY(cudaSetDevice(r->device));
    CUDA_TRY(cudaMemcpyAsync(host_image, r->image.p, pixel_count * sizeof(AtnVec4), cudaMemcpyDeviceToHost, r->stream));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    return ATN_OK;
}

ATN_API int atn_readback_aov(AtnRenderer* r, AtnVec4* normal_depth, AtnVec4* albedo_meshid, size_t pixel_count)
{
    if (!r || (!normal_depth && !albedo_meshid)) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_readback_aov: NULL argument");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_readback_aov: no scene uploaded");
    if (pixel_count != (size_t)r->width * r->height) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_readback_aov: pixel_count mismatch");
    CUDA_TRY(cudaSetDevice(r->device));
    if (normal_depth) CUDA_TRY(cudaMemcpyAsync(normal_depth, r->aov_nd.p, pixel_count * sizeof(AtnVec4), cudaMemcpyDeviceToHost, r->stream));
    if (albedo_meshid) CUDA_TRY(cudaMemcpyAsync(albedo_meshid, r->aov_am.p, pixel_count * sizeof(AtnVec4), cudaMemcpyDeviceToHost, r->stream));
    CUDA_TRY(cudaStreamSynchronize(r->stream));
    return ATN_OK;
}

ATN_API int atn_device_image(AtnRenderer* r, void** device_ptr, size_t* pixel_count)
{
    if (!r || !device_ptr) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_device_image: NULL argument");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_device_image: no scene uploaded");
    *device_ptr = r->image.p;
    if (pixel_count) *pixel_count = (size_t)r->width * r->height;
    return ATN_OK;
}

ATN_API int atn_get_stats(AtnRenderer* r, AtnStats* out)
{
    if (!r || !out) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_get_stats: NULL argument");
    const int rc = collect_stats(r);
    if (rc != ATN_OK) return rc;
    *out = r->stats;
    return ATN_OK;
}

static int query_launch(AtnRenderer* r, const void* rays_dev, const void* f_dev, size_t n, void* out_dev, bool any)
{
    if (n > (size_t)INT32_MAX - 64) return set_error(ATN_ERR_TOO_LARGE, "ray query: more than 2^31 rays in one call");
    cudaStream_t s = r->stream;
    CUDA_TRY(r->counters.resize(1));
    CUDA_TRY(cudaMemsetAsync(&r->counters.p->fetch_query, 0, sizeof(int32_t), s));
    CUDA_TRY(cudaMemsetAsync(&r->counters.p->n_exact_query, 0, sizeof(int32_t), s));
    if (r->exact_ties) CUDA_TRY(r->q_exact_query.resize(n));
    CUDA_TRY(r->visits.resize(1));
    CUDA_TRY(cudaMemsetAsync(r->visits.p, 0, sizeof(Visits), s));
    CUDA_TRY(cudaEventRecord(r->ev0, s));
    if (any) {
        QueryAnyPolicy pol{ (const AtnRay*)rays_dev, (const float*)f_dev, (uint8_t*)out_dev };
        launch_trace<QueryAnyPolicy>(r, s, pol, nullptr, (int)n, &r->counters.p->fetch_query, r->visits.p, r->q_exact_query.p, &r->counters.p->n_exact_query);
    }
    else {
        QueryClosestPolicy pol{ (const AtnRay*)rays_dev, (const float*)f_dev, (AtnHit*)out_dev };
        launch_trace<QueryClosestPolicy>(r, s, pol, nullptr, (int)n, &r->counters.p->fetch_query, r->visits.p, r->q_exact_query.p, &r->counters.p->n_exact_query);
    }
    r->stats.closest_rays = any ? 0 : n;
    r->stats.shadow_rays = any ? n : 0;
    return end_call(r, AtnRenderer::PEND_QUERY, 1, 0, 0);
}

ATN_API int atn_trace_closest_device(AtnRenderer* r, const void* rays_dev, const void* t_max_dev, size_t n, void* hits_dev)
{
    if (!r || !rays_dev || !hits_dev) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_trace_closest_device: NULL argument");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_trace_closest_device: no scene uploaded");
    if (n == 0) return ATN_OK;
    CUDA_TRY(cudaSetDevice(r->device));
    return query_launch(r, rays_dev, t_max_dev, n, hits_dev, false);
}

ATN_API int atn_trace_any_device(AtnRenderer* r, const void* rays_dev, const void* dist_dev, size_t n, void* visible_dev)
{
    if (!r || !rays_dev || !dist_dev || !visible_dev) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_trace_any_device: NULL argument");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_trace_any_device: no scene uploaded");
    if (n == 0) return ATN_OK;
    CUDA_TRY(cudaSetDevice(r->device));
    return query_launch(r, rays_dev, dist_dev, n, visible_dev, true);
}

ATN_API int atn_trace_closest(AtnRenderer* r, const AtnRay* rays, const float* t_max, size_t n, AtnHit* hits)
{
    if (!r || (n && (!rays || !hits))) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_trace_closest: NULL argument");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_trace_closest: no scene uploaded");
    if (n == 0) return ATN_OK;
    CUDA_TRY(cudaSetDevice(r->device));
    cudaStream_t s = r->stream;
    CUDA_TRY(r->q_rays.upload(rays, n, s));
    if (t_max) CUDA_TRY(r->q_f.upload(t_max, n, s));
    CUDA_TRY(r->q_hits.resize(n));
    int rc = query_launch(r, r->q_rays.p, t_max ? r->q_f.p : nullptr, n, r->q_hits.p, false);
    if (rc != ATN_OK) return rc;
    CUDA_TRY(cudaMemcpyAsync(hits, r->q_hits.p, n * sizeof(AtnHit), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return collect_stats(r);
}

ATN_API int atn_trace_any(AtnRenderer* r, const AtnRay* rays, const float* dist, size_t n, uint8_t* visible)
{
    if (!r || (n && (!rays || !dist || !visible))) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_trace_any: NULL argument");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_trace_any: no scene uploaded");
    if (n == 0) return ATN_OK;
    CUDA_TRY(cudaSetDevice(r->device));
    cudaStream_t s = r->stream;
    CUDA_TRY(r->q_rays.upload(rays, n, s));
    CUDA_TRY(r->q_f.upload(dist, n, s));
    CUDA_TRY(r->q_vis.resize(n));
    int rc = query_launch(r, r->q_rays.p, r->q_f.p, n, r->q_vis.p, true);
    if (rc != ATN_OK) return rc;
    CUDA_TRY(cudaMemcpyAsync(visible, r->q_vis.p, n, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return collect_stats(r);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// atn_trace_closest_static: closest-hit query that allocates nothing
// ---------------------------------------------------------------------------------------------
// atn_trace_closest stages the caller's host rays and hits in device buffers that it (re)allocates
// with cudaMalloc whenever a call is larger than the last one.  Some hosts cannot afford that: a
// renderer or simulation that was given a fixed device-memory budget at start-up and has spent it
// (an engine arena, a process that shares the GPU with a framework allocator which already took the
// rest), so cudaMalloc may simply fail mid-frame; and interactive code (picking, line-of-sight
// queries inside a frame loop) for which cudaMalloc's device-wide synchronisation is a visible
// stall.  For them the staging area lives in the library image itself: one file-scope __device__
// array per GPU, sized for kStaticMaxRays rays, present from the moment the library is loaded on a
// device until the process exits (about 4.2 MB per GPU).
constexpr size_t kStaticMaxRays = 65536;

struct StaticScratch {
    int32_t cursor;                    // work-fetch cursor of the persistent traversal kernel
    int32_t exact_count;               // rays handed to the exact walk (near ties)
    int32_t pad[62];                   // keeps the arrays below 256-byte aligned
    AtnRay rays[kStaticMaxRays];
    float t_max[kStaticMaxRays];
    AtnHit hits[kStaticMaxRays];
    int32_t exact_q[kStaticMaxRays];
};

// In host code this name is only a handle for cudaGetSymbolAddress; it is NOT a device address.
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

constexpr int kMaxDevices = 64;
static std::mutex g_static_mutex;
static StaticScratch* g_static_ptr[kMaxDevices] = {};     // device address of g_static_scratch, per GPU

// Looks up (once per GPU, with that GPU current) where g_static_scratch lives on the device.
static int static_scratch_for_device(int device, StaticScratch** out)
{
    if (device < 0 || device >= kMaxDevices)
        return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_trace_closest_static: device ordinal %d not supported (max %d)", device, kMaxDevices - 1);
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_ptr[device]) {
        void* p = nullptr;
        CUDA_TRY(cudaSetDevice(device));
        CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) return set_error(ATN_ERR_CUDA, "atn_trace_closest_static: cudaGetSymbolAddress returned NULL");
        g_static_ptr[device] = static_cast<StaticScratch*>(p);
    }
    *out = g_static_ptr[device];
    return ATN_OK;
}

// ---------------------------------------------------------------------------------------------
// Measurement aid (bench.py, scripts/l2_bandwidth.py): read bandwidth of a resident buffer.  One persistent launch walks a
// buffer of `bytes` (a power-of-two multiple of 32 B per thread step) `passes` times with 256-bit loads, so a buffer that
// fits the L2 is served from it for all but the first pass and the launch lasts milliseconds, not microseconds (a
// sequence of small copy launches measures launch latency instead, which is what the round-1 probe did below 32 MB).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_probe_read(const float4* __restrict__ buf, size_t n_chunks, int passes, float* sink)
{
    float acc = 0.0f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int p = 0; p < passes; p++) {
        // every pass a thread reads another part of the buffer, so that nothing is served from its SM's L1
        const size_t shift = ((size_t)p * 1048573u) % n_chunks;
        for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_chunks; i += stride) {
            size_t j = i + shift; if (j >= n_chunks) j -= n_chunks;
            const F4x2 v = ldg8(buf + 2 * j);
            acc += v.lo.x + v.hi.w;
        }
    }
    if (acc == 12345.678f) *sink = acc;          // keeps the loads alive
}

extern "C" ATN_API int atn_probe_read_bandwidth(int device_ordinal, size_t bytes, int passes, float* gb_per_s, float* ms_out)
{
    if (!gb_per_s || bytes < 4096 || passes < 1) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_probe_read_bandwidth: bad argument");
    CUDA_TRY(cudaSetDevice(device_ordinal));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device_ordinal));
    DevBuf<float4> buf; DevBuf<float> sink;
    const size_t n_chunks = bytes / 32;
    CUDA_TRY(buf.resize(n_chunks * 2)); CUDA_TRY(sink.resize(1));
    CUDA_TRY(cudaMemset(buf.p, 0, n_chunks * 32));
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0)); CUDA_TRY(cudaEventCreate(&e1));
    const int grid = prop.multiProcessorCount * 8;
    k_probe_read<<<grid, 256>>>(buf.p, n_chunks, 2, sink.p);                 // warm-up: the buffer is in the L2 afterwards
    CUDA_TRY(cudaEventRecord(e0));
    k_probe_read<<<grid, 256>>>(buf.p, n_chunks, passes, sink.p);
    CUDA_TRY(cudaEventRecord(e1));
    CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0.0f;
    CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *gb_per_s = (float)((double)n_chunks * 32.0 * passes / (ms * 1e-3) / 1e9);
    if (ms_out) *ms_out = ms;
    return ATN_OK;
}

extern "C" ATN_API int atn_trace_closest_static(AtnRenderer* r, const AtnRay* rays, const float* t_max, size_t n, AtnHit* hits)
{
    if (!r || (n && (!rays || !hits))) return set_error(ATN_ERR_INVALID_ARGUMENT, "atn_trace_closest_static: NULL argument");
    if (!r->has_scene) return set_error(ATN_ERR_NO_SCENE, "atn_trace_closest_static: no scene uploaded");
    if (n > kStaticMaxRays)
        return set_error(ATN_ERR_TOO_LARGE, "atn_trace_closest_static: %zu rays exceed the static staging area of %zu rays; "
                         "split the batch or use atn_trace_closest", n, kStaticMaxRays);
    if (n == 0) return ATN_OK;
    StaticScratch* sc = nullptr;
    const int rc = static_scratch_for_device(r->device, &sc);       // first call on a GPU: address look-up
    if (rc != ATN_OK) return rc;
    CUDA_TRY(cudaSetDevice(r->device));
    cudaStream_t s = r->stream;
    // The array keeps whatever the previous call left in it: the cursor must restart at zero; rays,
    // bounds and hits are fully overwritten for [0, n) and never read beyond n.
    CUDA_TRY(cudaMemsetAsync(&sc->cursor, 0, 2 * sizeof(int32_t), s));     // cursor and exact_count
    CUDA_TRY(cudaMemcpyAsync(sc->rays, rays, n * sizeof(AtnRay), cudaMemcpyHostToDevice, s));
    if (t_max) CUDA_TRY(cudaMemcpyAsync(sc->t_max, t_max, n * sizeof(float), cudaMemcpyHostToDevice, s));
    QueryClosestPolicy pol{ sc->rays, t_max ? sc->t_max : nullptr, sc->hits };
    CUDA_TRY(cudaEventRecord(r->ev0, s));
    launch_trace<QueryClosestPolicy>(r, s, pol, nullptr, (int)n, &sc->cursor, nullptr, sc->exact_q, &sc->exact_count);
    CUDA_TRY(cudaEventRecord(r->ev1, s));
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(hits, sc->hits, n * sizeof(AtnHit), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    // counters of this call (events and host fields allocate nothing); the visit counts are not taken on this path
    r->pending = AtnRenderer::PEND_NONE;
    r->stats = AtnStats{};
    r->stats.closest_rays = n;
    r->stats.kernel_launches = 1;
    float ms = 0.0f;
    CUDA_TRY(cudaEventElapsedTime(&ms, r->ev0, r->ev1));
    r->stats.device_ms = ms; r->stats.trace_ms = ms;
    return ATN_OK;
}
