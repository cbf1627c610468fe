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

#include "atn_trace.cuh"        // policies, k_trace, the exact walk
#include "atn_wavefront.cuh"    // generate / shade / gather, AO, SVGF, ReSTIR kernels

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
