// atn_wavefront.cuh -- the wavefront kernels of atn_kernels.cu around the traversal (one translation unit; included once,
// after atn_trace.cuh): k_generate, the three shade stages (k_surface, k_nee, k_bsdf) and the side pass k_side, the AO, SVGF and
// ReSTIR renderers' own kernels, k_gather.
#pragma once

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

