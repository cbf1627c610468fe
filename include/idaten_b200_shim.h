// idaten_b200_shim.h -- host-side C++ mirror of the reference's GPU renderer class over the C ABI.
//
// A maintainer of the reference swaps `idaten::PathTracing renderer_;` (src/device_renderer/main.cpp)
// for `idaten_b200::PathTracing renderer_;` and links libaten_b200.so: same method names, argument
// meaning and call order as src/libidaten/kernel/pathtracing.h:107-152 and
// src/libidaten/kernel/renderer.h:17-179.  This header is compiled against the reference's OWN
// headers ("aten.h"); it contains no CUDA and no copy of reference code -- it only reads the scene
// through the same aten::context getters Renderer::UpdateSceneData uses
// (src/libidaten/kernel/renderer.cpp:12-131) and hands the flat arrays to the C ABI.
//
// Differences from the reference, all deliberate:
//  * gltex is ignored (no graphics stack on the target): the image lives in a device float4 buffer,
//    read with ReadbackImage() or shared through DeviceImage().
//  * Q3: the reference's GPU upload rewrites ObjectParameter::mtx_id and mutates the host scene
//    (host_scene_context.cpp:285-307).  The shim emits the HOST indices (L2W at mtx_id, W2L at
//    mtx_id + 1) and does not touch the context.
//  * Q7: the frame counter starts at 0 like the host renderer (src/libaten/renderer/renderer.h:71).
//  * errors throw std::runtime_error instead of exit() (cudautil.h:9-54).
#pragma once

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <string>
#include <vector>

#include "aten.h"
#include "aten_b200.h"

namespace idaten_b200 {

// The structs must be layout-identical to the reference's, otherwise the zero-copy hand-over below
// would be wrong.  Checked here against the real headers.
static_assert(sizeof(AtnVec3) == sizeof(aten::vec3), "vec3");
static_assert(sizeof(AtnVec4) == sizeof(aten::vec4), "vec4");
static_assert(sizeof(AtnMat4) == sizeof(aten::mat4), "mat4");
static_assert(sizeof(AtnRay) == sizeof(aten::ray), "ray");
static_assert(sizeof(AtnHit) == sizeof(aten::Intersection), "Intersection");
static_assert(sizeof(AtnBvhNode) == sizeof(aten::GPUBvhNode), "GPUBvhNode");
static_assert(sizeof(AtnTriangleParameter) == sizeof(aten::TriangleParameter), "TriangleParameter");
static_assert(sizeof(AtnObjectParameter) == sizeof(aten::ObjectParameter), "ObjectParameter");
static_assert(sizeof(AtnMaterialParameter) == sizeof(aten::MaterialParameter), "MaterialParameter");
static_assert(sizeof(AtnLightParameter) == sizeof(aten::LightParameter), "LightParameter");
static_assert(sizeof(AtnCameraParameter) == sizeof(aten::CameraParameter), "CameraParameter");
static_assert(offsetof(AtnMaterialParameter, standard) == 44, "MaterialParameter::standard");
static_assert(offsetof(AtnMaterialParameter, albedoMap) == 28, "MaterialParameter::albedoMap");
static_assert(offsetof(AtnLightParameter, arealight_objid) == 68, "LightParameter::arealight_objid");
static_assert(offsetof(AtnObjectParameter, sphere_center) == 32, "ObjectParameter::sphere");

/// Owning flat copy of everything the path reads from aten::context -- the host half of
/// Renderer::UpdateSceneData (renderer.cpp:12-131).
struct FlatScene {
    std::vector<aten::ObjectParameter> objects;
    std::vector<aten::mat4> matrices;
    std::vector<aten::MaterialParameter> materials;
    std::vector<aten::LightParameter> lights;
    std::vector<aten::TriangleParameter> triangles;
    std::vector<aten::vec4> positions, normals;
    std::vector<std::vector<aten::GPUBvhNode>> nodes;
    std::vector<AtnTextureDesc> textures;
    std::vector<std::shared_ptr<aten::texture>> texture_refs;
    AtnSceneConfig config{};
    bool enable_alpha_blending{ false };

    // scratch for the blob view
    std::vector<const AtnBvhNode*> node_ptrs;
    std::vector<uint32_t> node_sizes;

    /// What Renderer::updateBVH reads (renderer.cpp:133-153): objects and matrices; the caller passes list 0 itself.
    /// Nothing that scales with the geometry is touched.
    void GatherTopLayer(const aten::context& ctxt)
    {
        const auto obj_num = ctxt.GetObjectNum();
        objects.clear();
        int32_t max_mtx = -1;
        for (size_t i = 0; i < obj_num; i++) {
            objects.push_back(ctxt.GetObject(static_cast<uint32_t>(i)));
            if (objects.back().type == aten::ObjectType::Instance) max_mtx = std::max(max_mtx, objects.back().mtx_id);
        }
        matrices.clear();
        for (int32_t i = 0; i <= max_mtx + 1 && max_mtx >= 0; i++) matrices.push_back(ctxt.GetMatrix(static_cast<uint32_t>(i)));
    }

    void Gather(const aten::context& ctxt, const std::vector<std::vector<aten::GPUBvhNode>>& node_lists)
    {
        const auto obj_num = ctxt.GetObjectNum();
        objects.clear();
        int32_t max_mtx = -1;
        for (size_t i = 0; i < obj_num; i++) {
            objects.push_back(ctxt.GetObject(static_cast<uint32_t>(i)));   // host mtx_id, unmodified
            if (objects.back().type == aten::ObjectType::Instance) {
                max_mtx = std::max(max_mtx, objects.back().mtx_id);
            }
        }
        matrices.clear();
        for (int32_t i = 0; i <= max_mtx + 1 && max_mtx >= 0; i++) {
            matrices.push_back(ctxt.GetMatrix(static_cast<uint32_t>(i)));
        }
        materials = ctxt.GetMetarialParemeters();
        lights = ctxt.GetLightParameters();
        triangles = ctxt.GetPrimitiveParameters();
        aten::tie(positions, normals) = ctxt.GetExtractedPosAndNmlInVertices();
        nodes = node_lists;

        textures.clear();
        texture_refs.clear();
        const auto tex_num = ctxt.GetTextureNum();
        for (int32_t i = 0; i < static_cast<int32_t>(tex_num); i++) {
            auto t = ctxt.GetTexture(i);
            texture_refs.push_back(t);
            AtnTextureDesc d{};
            d.texels = reinterpret_cast<const AtnVec4*>(t->colors().data());
            d.width = t->width();
            d.height = t->height();
            d.filter_mode = static_cast<int32_t>(t->GetFilterMode());
            d.address_mode = static_cast<int32_t>(t->GetAddressMode());
            textures.push_back(d);
        }

        const auto& cfg = ctxt.scene_rendering_config;
        config.bvh_hit_min = cfg.bvh_hit_min;
        config.bg_color = AtnVec3{ cfg.bg.bg_color.x, cfg.bg.bg_color.y, cfg.bg.bg_color.z };
        config.envmap_tex_idx = cfg.bg.envmap_tex_idx;
        config.avgIllum = cfg.bg.avgIllum;
        config.multiplyer = cfg.bg.multiplyer;
        config.enable_env_map = cfg.bg.enable_env_map ? 1 : 0;
        enable_alpha_blending = cfg.enable_alpha_blending;
        const auto& bb = ctxt.GetSceneBoundingBox();
        const auto bmin = bb.minPos();
        const auto bmax = bb.maxPos();
        config.scene_bbox_min[0] = bmin.x; config.scene_bbox_min[1] = bmin.y; config.scene_bbox_min[2] = bmin.z;
        config.scene_bbox_max[0] = bmax.x; config.scene_bbox_max[1] = bmax.y; config.scene_bbox_max[2] = bmax.z;
    }

    AtnSceneBlob View()
    {
        node_ptrs.clear();
        node_sizes.clear();
        for (const auto& l : nodes) {
            node_ptrs.push_back(reinterpret_cast<const AtnBvhNode*>(l.data()));
            node_sizes.push_back(static_cast<uint32_t>(l.size()));
        }
        AtnSceneBlob b{};
        b.objects = reinterpret_cast<const AtnObjectParameter*>(objects.data());
        b.object_count = static_cast<uint32_t>(objects.size());
        b.matrices = reinterpret_cast<const AtnMat4*>(matrices.data());
        b.matrix_count = static_cast<uint32_t>(matrices.size());
        b.materials = reinterpret_cast<const AtnMaterialParameter*>(materials.data());
        b.material_count = static_cast<uint32_t>(materials.size());
        b.lights = reinterpret_cast<const AtnLightParameter*>(lights.data());
        b.light_count = static_cast<uint32_t>(lights.size());
        b.triangles = reinterpret_cast<const AtnTriangleParameter*>(triangles.data());
        b.triangle_count = static_cast<uint32_t>(triangles.size());
        b.positions = reinterpret_cast<const AtnVec4*>(positions.data());
        b.normals = reinterpret_cast<const AtnVec4*>(normals.data());
        b.vertex_count = static_cast<uint32_t>(positions.size());
        b.node_lists = node_ptrs.data();
        b.node_list_sizes = node_sizes.data();
        b.node_list_count = static_cast<uint32_t>(nodes.size());
        b.textures = textures.data();
        b.texture_count = static_cast<uint32_t>(textures.size());
        b.config = config;
        b.enable_alpha_blending = enable_alpha_blending ? 1 : 0;
        return b;
    }

    /// Serialise the blob (+ camera, size, seeds) as a `.atnscene` file: the portable form of a
    /// scene for machines that have no reference checkout (tests/golden, the GPU box).
    /// Layout: see aten_b200/scene_io.py.
    bool Save(const std::string& path, const aten::CameraParameter& camera, int32_t width, int32_t height,
              const std::vector<uint32_t>& seeds)
    {
        FILE* fp = fopen(path.c_str(), "wb");
        if (!fp) return false;
        auto w = [&](const void* p, size_t n) { if (n) fwrite(p, 1, n, fp); };
        const char magic[8] = { 'A','T','N','S','C','N','0','1' };
        w(magic, 8);
        uint32_t counts[10] = {
            (uint32_t)objects.size(), (uint32_t)matrices.size(), (uint32_t)materials.size(),
            (uint32_t)lights.size(), (uint32_t)triangles.size(), (uint32_t)positions.size(),
            (uint32_t)nodes.size(), (uint32_t)textures.size(), (uint32_t)seeds.size(),
            enable_alpha_blending ? 1u : 0u };      // flags: bit 0 = enable_alpha_blending
        w(counts, sizeof(counts));
        w(&config, sizeof(config));
        w(&camera, sizeof(camera));
        int32_t wh[2] = { width, height };
        w(wh, sizeof(wh));
        w(objects.data(), objects.size() * sizeof(aten::ObjectParameter));
        w(matrices.data(), matrices.size() * sizeof(aten::mat4));
        w(materials.data(), materials.size() * sizeof(aten::MaterialParameter));
        w(lights.data(), lights.size() * sizeof(aten::LightParameter));
        w(triangles.data(), triangles.size() * sizeof(aten::TriangleParameter));
        w(positions.data(), positions.size() * sizeof(aten::vec4));
        w(normals.data(), normals.size() * sizeof(aten::vec4));
        for (const auto& l : nodes) { uint32_t n = (uint32_t)l.size(); w(&n, 4); }
        for (const auto& l : nodes) w(l.data(), l.size() * sizeof(aten::GPUBvhNode));
        for (const auto& t : textures) {
            int32_t hdr[4] = { t.width, t.height, t.filter_mode, t.address_mode };
            w(hdr, sizeof(hdr));
            w(t.texels, sizeof(AtnVec4) * (size_t)t.width * (size_t)t.height);
        }
        w(seeds.data(), seeds.size() * 4);
        fclose(fp);
        return true;
    }
};

/// Same public surface as idaten::PathTracing (pathtracing.h:107-152) + the idaten::Renderer
/// methods the apps call (renderer.h:61-116).
class PathTracing {
public:
    enum class Mode { PT, AOV };

    PathTracing() = default;
    explicit PathTracing(int device_ordinal) : device_(device_ordinal) {}
    ~PathTracing() { if (h_) atn_destroy(h_); }
    PathTracing(const PathTracing&) = delete;
    PathTracing& operator=(const PathTracing&) = delete;

    /// idaten::PathTracing::UpdateSceneData (pathtracing.h:117-125).  `gltex` is ignored.
    void UpdateSceneData(
        uint32_t /*gltex*/,
        int32_t width, int32_t height,
        const aten::CameraParameter& camera,
        const aten::context& scene_ctxt,
        const std::vector<std::vector<aten::GPUBvhNode>>& nodes,
        uint32_t /*advance_prim_num*/ = 0,
        uint32_t /*advance_vtx_num*/ = 0,
        std::function<const aten::Grid*(const aten::context&)> /*proxy*/ = nullptr)
    {
        Ensure();
        flat_.Gather(scene_ctxt, nodes);
        auto blob = flat_.View();
        if (encoded_cache_path_.empty())
            Check(atn_upload_scene(h_, &blob, reinterpret_cast<const AtnCameraParameter*>(&camera), width, height));
        else {
            int hit = 0;
            Check(atn_upload_scene_cached(h_, &blob, reinterpret_cast<const AtnCameraParameter*>(&camera), width, height,
                                          encoded_cache_path_.c_str(), &hit));
            encoded_cache_hit_ = hit != 0;
        }
        // Renderer::initSamplerParameter (renderer.h:113-124)
        const auto& rnd = aten::getRandom();
        Check(atn_set_seeds(h_, rnd.data(), static_cast<uint32_t>(rnd.size())));
    }

    /// No counterpart in the reference.  false (default): textures are read with the reference's HOST rule
    /// (aten::texture::at), which is what its CPU path tracer -- the parity oracle -- computes.  true: with the rule of the
    /// renderer this class stands in for, tex2DLod with each texture's filter / address mode (sample_texture.h:18-35).
    void SetDeviceTextureFetch(bool device_rule) { Ensure(); Check(atn_set_texture_fetch(h_, device_rule ? 1 : 0)); }

    /// No counterpart in the reference: the next UpdateSceneData keeps the ENCODED bottom trees in this file and skips
    /// the encoding step when the file matches the scene (atn_upload_scene_cached); "" switches it off.
    void SetEncodedTreeCache(const std::string& path) { encoded_cache_path_ = path; }
    bool EncodedTreeCacheHit() const { return encoded_cache_hit_; }

    /// No counterpart in the reference: upper bound on the paths of one wavefront, i.e. on the device memory of the path
    /// state (~265 B per path; default 32 Mi paths).  Before UpdateSceneData (atn_set_wavefront_paths).
    void SetWavefrontPaths(uint64_t max_paths) { Ensure(); Check(atn_set_wavefront_paths(h_, max_paths)); }

    /// sbvh::importTree (src/libaten/accelerator/sbvh.cpp:1345-1436) without an aten::sbvh object: the node list of a
    /// `.sbvh` file, ready to stand as one bottom-layer list of the `nodes` argument of UpdateSceneData.  Voxel material
    /// ids are re-mapped through `scene_ctxt`'s material names, as importTree does with FindMaterialIdxByName.
    static std::vector<aten::GPUBvhNode> ImportSbvh(const std::string& path, int32_t offset_tri_idx = 0,
                                                    const std::vector<std::string>* material_names = nullptr)
    {
        std::vector<const char*> names;
        if (material_names) for (const auto& n : *material_names) names.push_back(n.c_str());
        AtnBvhNode* nodes = nullptr;
        uint32_t count = 0;
        Check(atn_sbvh_read(path.c_str(), offset_tri_idx, material_names ? names.data() : nullptr, (uint32_t)names.size(),
                            &nodes, &count, nullptr));
        static_assert(sizeof(aten::GPUBvhNode) == sizeof(AtnBvhNode), "node layout");
        std::vector<aten::GPUBvhNode> out(count);
        if (count) memcpy(out.data(), nodes, count * sizeof(AtnBvhNode));
        atn_free(nodes);
        return out;
    }

    /// PathTracingImplBase::render (pathtracing.cpp:49-89)
    void render(int32_t width, int32_t height, int32_t maxSamples, int32_t maxBounce)
    {
        Check(atn_render(h_, width, height, maxSamples, maxBounce, /*rrBounce (pathtracing.cpp:97)*/ 3));
    }

    void updateCamera(const aten::CameraParameter& camera)
    {
        Check(atn_update_camera(h_, reinterpret_cast<const AtnCameraParameter*>(&camera)));
    }

    /// Renderer::updateBVH (renderer.cpp:133-153)
    void updateBVH(const aten::context& scene_ctxt, const std::vector<std::vector<aten::GPUBvhNode>>& nodes)
    {
        if (nodes.empty()) return;
        flat_.GatherTopLayer(scene_ctxt);
        Check(atn_update_top_layer(
            h_, reinterpret_cast<const AtnBvhNode*>(nodes[0].data()), (uint32_t)nodes[0].size(),
            reinterpret_cast<const AtnObjectParameter*>(flat_.objects.data()), (uint32_t)flat_.objects.size(),
            reinterpret_cast<const AtnMat4*>(flat_.matrices.data()), (uint32_t)flat_.matrices.size()));
    }

    void updateMaterial(const std::vector<aten::MaterialParameter>& mtrls)
    {
        Check(atn_update_materials(h_, reinterpret_cast<const AtnMaterialParameter*>(mtrls.data()), (uint32_t)mtrls.size()));
    }

    void updateLight(const aten::context& scene_ctxt, bool /*is_npr_target_light*/ = false)
    {
        const auto lights = scene_ctxt.GetLightParameters();
        Check(atn_update_lights(h_, reinterpret_cast<const AtnLightParameter*>(lights.data()), (uint32_t)lights.size()));
    }

    /// Renderer::UpdateTexture (renderer.cpp:236-247)
    void UpdateTexture(int32_t idx, const aten::context& ctxt)
    {
        auto tex = ctxt.GetTexture(idx);
        AtnTextureDesc d{};
        d.texels = reinterpret_cast<const AtnVec4*>(tex->colors().data());
        d.width = static_cast<int32_t>(tex->width()); d.height = static_cast<int32_t>(tex->height());
        Check(atn_update_texture(h_, static_cast<uint32_t>(idx), &d));
    }

    void UpdateSceneRenderingConfig(const aten::context& ctxt)
    {
        FlatScene tmp;
        tmp.config = flat_.config;
        const auto& cfg = ctxt.scene_rendering_config;
        tmp.config.bvh_hit_min = cfg.bvh_hit_min;
        tmp.config.bg_color = AtnVec3{ cfg.bg.bg_color.x, cfg.bg.bg_color.y, cfg.bg.bg_color.z };
        tmp.config.envmap_tex_idx = cfg.bg.envmap_tex_idx;
        tmp.config.avgIllum = cfg.bg.avgIllum;
        tmp.config.multiplyer = cfg.bg.multiplyer;
        tmp.config.enable_env_map = cfg.bg.enable_env_map ? 1 : 0;
        flat_.config = tmp.config;
        flat_.enable_alpha_blending = cfg.enable_alpha_blending;
        Check(atn_update_config(h_, &flat_.config));
        Check(atn_set_alpha_blending(h_, cfg.enable_alpha_blending ? 1 : 0));
    }

    void reset() { Check(atn_reset(h_)); }
    uint32_t GetFrameCount() const { uint32_t f = 0; atn_get_frame(h_, &f); return f; }
    void SetEnableProgressive(bool b) { progressive_ = b; Ensure(); Check(atn_set_progressive(h_, b ? 1 : 0)); }
    /// No counterpart in the reference: ATN_GATHER_SUM keeps sum + count in the image (the multi-GPU reduce form).
    void SetGatherMode(AtnGatherMode m) { Ensure(); progressive_ = (m == ATN_GATHER_PROGRESSIVE); Check(atn_set_gather_mode(h_, (int)m)); }
    bool IsEnableProgressive() const { return progressive_; }
    void setHitDistanceLimit(float d) { Check(atn_set_hit_distance_limit(h_, d)); }
    void SetRenderingMode(Mode m) { mode_ = m; }
    Mode GetRenderingMode() const { return mode_; }
    void* GetCudaStream() const { void* s = nullptr; atn_get_stream(h_, &s); return s; }

    /// Replacement for the GL surface: copy the RGBA32F image to the host.
    void ReadbackImage(aten::vec4* dst, size_t pixel_count)
    {
        Check(atn_readback(h_, reinterpret_cast<AtnVec4*>(dst), pixel_count));
    }
    void* DeviceImage() { void* p = nullptr; size_t n = 0; Check(atn_device_image(h_, &p, &n)); return p; }

    AtnRenderer* handle() { Ensure(); return h_; }
    FlatScene& flat() { return flat_; }

protected:
    void Ensure()
    {
        if (h_) return;
        Check(atn_create(device_, &h_));
        // idaten::PathTracing starts with m_enableProgressive{false} (pathtracing.h:103): every render() overwrites the image
        Check(atn_set_progressive(h_, progressive_ ? 1 : 0));
    }
    static void Check(int rc) { if (rc != ATN_OK) throw std::runtime_error(atn_last_error()); }

    AtnRenderer* h_{ nullptr };
    int device_{ 0 };
    bool progressive_{ false };
    std::string encoded_cache_path_;
    bool encoded_cache_hit_{ false };
    Mode mode_{ Mode::PT };
    FlatScene flat_;
};

/// idaten::AORenderer (src/libidaten/ao/ao.h:9-52): PathTracing whose render() is the ambient-occlusion pass.
class AORenderer : public PathTracing {
public:
    using PathTracing::PathTracing;
    int32_t getNumRays() const { return num_rays_; }
    void setNumRays(int32_t num) { num_rays_ = num; }
    float getRadius() const { return radius_; }
    void setRadius(float radius) { radius_ = radius; }

    /// AORenderer::OnRender (src/libidaten/ao/ao.cpp:9-51); maxBounce is ignored there as well
    void render(int32_t width, int32_t height, int32_t maxSamples, int32_t /*maxBounce*/)
    {
        Check(atn_render_ao(h_, width, height, maxSamples, num_rays_, radius_));
    }

private:
    int32_t num_rays_{ 1 };     // m_ao_num_rays
    float radius_{ 1.0f };      // m_ao_radius
};

/// idaten::SVGFPathTracing (src/libidaten/svgf/svgf.h:13-215): one render() call = one frame of the SVGF sequence.
/// The GL G-buffer / motion-depth textures of SetGBuffer() are replaced by a host buffer of motion vectors.
class SVGFPathTracing : public PathTracing {
public:
    enum Mode { SVGF, TF, PT, VAR, AOVar };
    using PathTracing::PathTracing;

    Mode GetMode() const { return mode_; }
    void SetMode(Mode mode)
    {
        if (mode != SVGF && mode != PT) throw std::runtime_error("idaten_b200::SVGFPathTracing: only Mode::SVGF and Mode::PT are implemented");
        const Mode prev = mode_;
        mode_ = mode;
        if (prev != mode_) reset();
    }

    /// SVGFPathTracing::render (svgf.cpp:28-125).  maxSamples is 1 in every caller of the reference; more than one
    /// sample per frame is not implemented (the host twin's sample loop stops at the first terminated path anyway,
    /// src/libaten/renderer/svgf/svgf.cpp:494-519).
    void render(int32_t width, int32_t height, int32_t maxSamples, int32_t maxBounce)
    {
        if (mode_ == PT) { PathTracing::render(width, height, maxSamples, maxBounce); return; }
        if (maxSamples != 1) throw std::runtime_error("idaten_b200::SVGFPathTracing::render: maxSamples must be 1");
        Check(atn_render_svgf(h_, width, height, maxBounce, /*rrBounce*/ 3, atrous_iters_));
    }

    /// Stands in for SetGBuffer(gltexGbuffer, gltexMotionDepthbuffer): W*H float4, .xy = motion in screen fractions.
    void SetMotionDepthBuffer(const aten::vec4* motion_depth, size_t pixel_count)
    {
        Check(atn_set_motion_depth(h_, reinterpret_cast<const AtnVec4*>(motion_depth), pixel_count));
    }

    void SetAtrousIterCount(int32_t n) { atrous_iters_ = n; }      // SVGFParams::atrous_iter_cnt (svgf_types.h:74)

    /// The sequence restarts without history.  (idaten::Renderer::reset() sets the frame counter to 1, which makes the
    /// reference re-use whatever history its buffers hold; here frame 0 = no history.)
    void reset() { PathTracing::reset(); }

private:
    Mode mode_{ SVGF };
    int32_t atrous_iters_{ 5 };
};

/// idaten::ReSTIRPathTracing (src/libidaten/restir/restir.h): one render() call = one frame of the ReSTIR sequence.
class ReSTIRPathTracing : public PathTracing {
public:
    enum class Mode { ReSTIR, PT, AOVar };
    using PathTracing::PathTracing;

    Mode GetMode() const { return mode_; }
    void SetMode(Mode mode)
    {
        if (mode == Mode::AOVar) throw std::runtime_error("idaten_b200::ReSTIRPathTracing: Mode::AOVar (debug display) is not implemented");
        const Mode prev = mode_;
        mode_ = mode;
        if (prev != mode_) reset();
    }

    /// ReSTIRPathTracing::OnRender (restir.cpp:47-123); one sample per frame, as every caller of the reference passes
    void render(int32_t width, int32_t height, int32_t maxSamples, int32_t maxBounce)
    {
        if (mode_ == Mode::PT) { PathTracing::render(width, height, maxSamples, maxBounce); return; }
        if (maxSamples != 1) throw std::runtime_error("idaten_b200::ReSTIRPathTracing::render: maxSamples must be 1");
        Check(atn_render_restir(h_, width, height, maxBounce, /*rrBounce (restir.cpp:53)*/ 3));
    }

    /// Stands in for SetGBuffer(gltexGbuffer, gltexMotionDepthbuffer): W*H float4, .xy = motion in screen fractions.
    void SetMotionDepthBuffer(const aten::vec4* motion_depth, size_t pixel_count)
    {
        Check(atn_set_motion_depth(h_, reinterpret_cast<const AtnVec4*>(motion_depth), pixel_count));
    }

private:
    Mode mode_{ Mode::ReSTIR };
};

}  // namespace idaten_b200
