/*
 * aten_b200 -- C ABI of the B200-native path-tracing hot path.
 *
 * This is the drop-in boundary for the reference's GPU renderer object
 * (idaten::PathTracing : PathTracingImplBase : Renderer, reference
 * src/libidaten/kernel/pathtracing.h:13-152 and src/libidaten/kernel/renderer.h:17-179).
 * Every entry point below names the reference interface it replaces.  All arguments are plain
 * pointers and sizes; no CUDA, torch or C++ types cross the boundary.  The POD structs mirror the
 * reference's own parameter structs byte for byte (sizes/offsets are static_assert'ed in
 * aten_b200/csrc and again against the real reference headers in include/idaten_b200_shim.h), so
 * the arrays returned by aten::context's getters can be passed without conversion.
 *
 * Threading: one handle owns one CUDA device + one stream; a handle is not re-entrant (the
 * reference renderer is single threaded on stream 0, renderer.h:161).  Different handles may be
 * used from different threads.
 *
 * Errors: every function returns ATN_OK (0) or a negative AtnStatus; atn_last_error() returns the
 * message of the calling thread's last failure.  Unlike the reference (cudautil.h:9-54, which
 * prints, cudaDeviceReset()s and exit()s) the library never terminates the process.  There is no
 * CPU fallback: without a CUDA device every call fails with ATN_ERR_CUDA.
 */
#ifndef ATEN_B200_H_
#define ATEN_B200_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define ATN_API __declspec(dllexport)
#else
#define ATN_API __attribute__((visibility("default")))
#endif

typedef enum AtnStatus {
    ATN_OK = 0,
    ATN_ERR_INVALID_ARGUMENT = -1,
    ATN_ERR_CUDA = -2,
    ATN_ERR_NO_SCENE = -3,
    ATN_ERR_UNSUPPORTED = -4,
    ATN_ERR_TOO_LARGE = -5,
} AtnStatus;

/* ---- POD mirrors of the reference's parameter structs ---------------- */

typedef struct AtnVec3 { float x, y, z; } AtnVec3;                 /* aten::vec3, 12 B */
typedef struct AtnVec4 { float x, y, z, w; } AtnVec4;              /* aten::vec4, 16 B */
typedef struct AtnMat4 { float m[16]; } AtnMat4;                   /* aten::mat4, row major, 64 B */

/* aten::ray (src/libaten/math/ray.h:7-83), 24 B */
typedef struct AtnRay { AtnVec3 org; AtnVec3 dir; } AtnRay;

/* aten::Intersection (src/libaten/scene/hit_parameter.h:28-64), 32 B */
typedef struct AtnHit {
    float t;          /* AT_MATH_INF (FLT_MAX) on miss */
    int32_t objid;    /* -1 on miss */
    int32_t mtrlid;
    int32_t meshid;
    int32_t primid;   /* hit.tri.id */
    float a, b;       /* barycentrics, p = (1-a-b) v0 + a v1 + b v2 */
    int32_t isVoxel;
} AtnHit;

/* aten::ThreadedBvhNode == ThreadedSbvhNode == GPUBvhNode
 * (src/libaten/accelerator/threaded_bvh.h:13-56, sbvh.h:19-66), 48 B; all links/ids are floats. */
typedef struct AtnBvhNode {
    float boxmin[3]; float hit;
    float boxmax[3]; float miss;
    float object_id;  /* sbvh: isleaf */
    float primid;     /* sbvh: triid */
    float exid;       /* bit field mainExid:15 lodExid:15 hasLod:1 noExternal:1; sbvh: voxeldepth */
    float meshid;     /* sbvh: mtrlid */
} AtnBvhNode;

/* aten::TriangleParameter (src/libaten/geometry/geomparam.h:66-98), 32 B */
typedef struct AtnTriangleParameter {
    int32_t idx[3]; float padding;
    float area; int32_t needNormal; int32_t mtrlid; int32_t mesh_id;
} AtnTriangleParameter;

/* aten::ObjectParameter (geomparam.h:23-64), 64 B */
typedef struct AtnObjectParameter {
    int32_t type;         /* 0 Polygons, 1 Instance, 2 Sphere */
    float area;
    int32_t object_id;    /* Instance: index of the real object */
    int32_t mtx_id;       /* HOST semantics: absolute index of L2W; W2L is mtx_id + 1 */
    int32_t triangle_id;
    int32_t triangle_num;
    int32_t light_id;
    int32_t pad0;
    float sphere_center[3]; float sphere_radius;
    int32_t sphere_mtrl_id; int32_t pad1[3];
} AtnObjectParameter;

/* aten::StandardMaterialParameter (src/libaten/material/material.h:66-120), 48 B */
typedef struct AtnStandardMaterialParameter {
    float ior, roughness, shininess, subsurface, metallic, specular, specularTint, anisotropic,
          sheen, sheenTint, clearcoat, clearcoatGloss;
} AtnStandardMaterialParameter;

/* aten::MaterialParameter (material.h:242-325), 248 B */
typedef struct AtnMaterialParameter {
    AtnVec4 baseColor;                 /*   0 */
    int32_t type;                      /*  16  aten::MaterialType */
    uint32_t attrib;                   /*  20  bit0 emissive, bit1 singular, bit2 translucent, bit3 glossy */
    uint16_t id;                       /*  24 */
    uint8_t isIdealRefraction;         /*  26 */
    uint8_t is_medium;                 /*  27 */
    int32_t albedoMap;                 /*  28 */
    int32_t normalMap;                 /*  32 */
    int32_t roughnessMap;              /*  36 */
    int32_t stencil_type;              /*  40 */
    AtnStandardMaterialParameter standard; /* 44 (union with carpaint, 64 B) */
    uint8_t carpaint_tail[16];         /*  92 */
    uint8_t medium[32];                /* 108 */
    uint8_t toon[100];                 /* 140 */
    uint8_t feature_line[8];           /* 240 */
} AtnMaterialParameter;

enum {  /* aten::MaterialType (material.h:41-64) */
    ATN_MTRL_EMISSIVE = 0, ATN_MTRL_DIFFUSE = 1, ATN_MTRL_ORENNAYAR = 2, ATN_MTRL_SPECULAR = 3,
    ATN_MTRL_REFRACTION = 4, ATN_MTRL_GGX = 5, ATN_MTRL_BECKMAN = 6, ATN_MTRL_VELVET = 7,
    ATN_MTRL_MICROFACET_REFRACTION = 8, ATN_MTRL_RETROREFLECTIVE = 9, ATN_MTRL_CARPAINT = 10,
    ATN_MTRL_DISNEY = 11, ATN_MTRL_TOON = 12, ATN_MTRL_STYLIZED = 13
};

/* aten::LightParameter (src/libaten/light/light_parameter.h:47-86), 80 B */
typedef struct AtnLightParameter {
    AtnVec4 pos; AtnVec4 dir;
    int32_t type;            /* 0 Area, 1 IBL, 2 Direction, 3 Point, 4 Spot */
    AtnVec3 light_color;
    float innerAngle, outerAngle;
    uint32_t attrib;         /* bit0 singular, bit1 infinite, bit2 IBL */
    float scale;
    float intensity;
    int32_t arealight_objid;
    int32_t envmapidx;
    int32_t padding;
} AtnLightParameter;

/* aten::CameraParameter (src/libaten/camera/camera.h:15-38), 124 B */
typedef struct AtnCameraParameter {
    AtnVec3 origin, lookat; float aspect; AtnVec3 center, u, v, dir, right, up;
    float dist, vfov; int32_t width, height; float znear, zfar;
} AtnCameraParameter;

/* The fields of aten::SceneRenderingConfig the path reads (scene_rendering_config.h:7-41). */
typedef struct AtnSceneConfig {
    float bvh_hit_min;            /* <= 0: use AT_MATH_EPSILON */
    AtnVec3 bg_color;             /* BackgroundResource */
    int32_t envmap_tex_idx;
    float avgIllum;
    float multiplyer;
    int32_t enable_env_map;
    float scene_bbox_min[3];      /* aten::context::GetSceneBoundingBox() (IBL light distance) */
    float scene_bbox_max[3];
} AtnSceneConfig;

/* aten::texture: float4 texels, row major (src/libaten/image/texture.h).  By default the path point-samples with the
 * host rule int(u*(w-1)) + repeat wrap (aten::texture::at, texture.cpp:64-75), as the CPU oracle does; with
 * atn_set_texture_fetch(1) every texture is fetched the way the reference's GPU renderer fetches it, with its own
 * filter and address mode. */
typedef struct AtnTextureDesc {
    const AtnVec4* texels;
    int32_t width, height;
    int32_t filter_mode;   /* aten::TextureFilterMode: 0 Point, 1 Linear */
    int32_t address_mode;  /* aten::TextureAddressMode: 0 Wrap, 1 Clamp, 2 Mirror, 3 Border */
} AtnTextureDesc;

/* Everything Renderer::UpdateSceneData reads from aten::context
 * (src/libidaten/kernel/renderer.cpp:12-131). The library copies all of it; the caller keeps
 * ownership. */
typedef struct AtnSceneBlob {
    const AtnObjectParameter* objects;       uint32_t object_count;
    const AtnMat4* matrices;                 uint32_t matrix_count;
    const AtnMaterialParameter* materials;   uint32_t material_count;
    const AtnLightParameter* lights;         uint32_t light_count;
    const AtnTriangleParameter* triangles;   uint32_t triangle_count;
    const AtnVec4* positions;                /* (x,y,z,u) per vertex */
    const AtnVec4* normals;                  /* (nx,ny,nz,v) per vertex */
    uint32_t vertex_count;
    const AtnBvhNode* const* node_lists;     /* list 0 = top layer, list k = bottom layer k */
    const uint32_t* node_list_sizes;
    uint32_t node_list_count;
    const AtnTextureDesc* textures;          uint32_t texture_count;
    AtnSceneConfig config;
    int32_t enable_alpha_blending;           /* SceneRenderingConfig::enable_alpha_blending (scene_rendering_config.h:21) */
} AtnSceneBlob;

/* Work counters of the last atn_render / atn_trace_* call. */
typedef struct AtnStats {
    uint64_t closest_rays;     /* closest-hit traversals launched (primary + bounce rays) */
    uint64_t shadow_rays;      /* shadow-ray traversals launched */
    uint64_t camera_samples;   /* W*H*spp of the call */
    uint64_t kernel_launches;  /* CUDA kernels launched by the call */
    float    device_ms;        /* CUDA-event time of the call on the handle's stream */
    float    trace_ms;         /* sum of CUDA-event times of the traversal kernels (if profiling on) */
    uint64_t nodes_visited;    /* 128-byte QuadNodes fetched by the traversal kernels (renders and ray queries); counted only
                                  with atn_set_profiling(2), 0 otherwise */
    uint64_t tris_tested;      /* triangle tests run by the traversal kernels; likewise */
    uint64_t exact_rays;       /* rays whose hit depended on a near tie and was decided by the reference-order walk */
} AtnStats;

typedef struct AtnRenderer AtnRenderer;

/* ---- lifetime -------------------------------------------------------------------------------- */
/* idaten::Renderer::Renderer() (renderer.cpp:5-9) */
ATN_API int atn_create(int device_ordinal, AtnRenderer** out);
ATN_API int atn_destroy(AtnRenderer* r);
ATN_API const char* atn_last_error(void);
ATN_API const char* atn_version(void);

/* ---- scene upload ----------------------------------------------------------------------------- */
/* Renderer::UpdateSceneData (renderer.cpp:12-131 / pathtracing.h:117-125): copies the scene,
 * re-encodes the threaded node lists into the traversal layout and allocates the
 * W*H path state. */
ATN_API int atn_upload_scene(AtnRenderer* r, const AtnSceneBlob* scene,
                             const AtnCameraParameter* camera, int32_t width, int32_t height);
/* The same upload for a scene whose bottom trees are large: `cache_path` names a file that holds the
 * ENCODED bottom trees (the 4-slot nodes and gathered triangle records the node lists are turned into).  When it matches
 * exactly these node lists, triangles and vertex positions (hash) and this library's encoder version, the encoding step
 * is skipped (*cache_hit = 1); otherwise the scene is encoded as usual and the file is written for the next time
 * (*cache_hit = 0).  The top layer is never cached.  cache_hit may be NULL. */
ATN_API int atn_upload_scene_cached(AtnRenderer* r, const AtnSceneBlob* scene,
                                    const AtnCameraParameter* camera, int32_t width, int32_t height,
                                    const char* cache_path, int* cache_hit);

/* ---- .sbvh tree files -------------------------------------------------------------------------- */
typedef struct AtnSbvhInfo {
    uint8_t version[4];
    uint32_t max_depth;
    uint32_t voxel_material_count;
    float boxmin[3], boxmax[3];
} AtnSbvhInfo;
/* sbvh::importTree(ctxt, path, offsetTriIdx) (src/libaten/accelerator/sbvh.cpp:1345-1436): reads a tree file written by
 * the reference (or by atn_sbvh_write) into a node array usable as one bottom-layer list of AtnSceneBlob.  Triangle
 * numbers get offset_tri_idx added; voxel material ids are re-mapped through the file's name table to indices into
 * material_names (context::FindMaterialIdxByName; pass NULL to keep the file's ids).  *nodes is malloc'ed: release it
 * with atn_free.  info may be NULL. */
ATN_API int atn_sbvh_read(const char* path, int32_t offset_tri_idx, const char* const* material_names, uint32_t material_count,
                          AtnBvhNode** nodes, uint32_t* node_count, AtnSbvhInfo* info);
/* sbvh::exportTree's file layout (sbvh.cpp:1236-1340) for a node list the caller holds. */
ATN_API int atn_sbvh_write(const char* path, const AtnBvhNode* nodes, uint32_t node_count, uint32_t max_depth,
                           const float boxmin[3], const float boxmax[3],
                           const int32_t* voxel_material_ids, const char* const* voxel_material_names, uint32_t voxel_material_count);
ATN_API void atn_free(void* p);

/* Renderer::updateCamera (renderer.h:85-88) */
ATN_API int atn_update_camera(AtnRenderer* r, const AtnCameraParameter* camera);
/* Renderer::updateBVH (renderer.cpp:133-153): new top layer (list 0), objects and matrices. */
ATN_API int atn_update_top_layer(AtnRenderer* r, const AtnBvhNode* top_nodes, uint32_t node_count,
                                 const AtnObjectParameter* objects, uint32_t object_count,
                                 const AtnMat4* matrices, uint32_t matrix_count);
/* Renderer::updateMaterial (renderer.cpp:190-200) */
ATN_API int atn_update_materials(AtnRenderer* r, const AtnMaterialParameter* m, uint32_t count);
/* Renderer::UpdateTexture(idx, ctxt) (src/libidaten/kernel/renderer.cpp:236-247): new texels for texture `index`;
 * the reference copies into the existing CUDA texture resource, so the size stays what it was at upload. */
ATN_API int atn_update_texture(AtnRenderer* r, uint32_t index, const AtnTextureDesc* texture);
/* Renderer::updateLight (renderer.cpp:202-220) */
ATN_API int atn_update_lights(AtnRenderer* r, const AtnLightParameter* l, uint32_t count);
/* Renderer::UpdateSceneRenderingConfig (renderer.h:103-106) */
ATN_API int atn_update_config(AtnRenderer* r, const AtnSceneConfig* cfg);
/* Which texel a texture lookup returns.  0 (default): the reference's HOST rule, aten::texture::at -- point
 * sample at int(u * (w - 1)) with repeat wrap (src/libaten/image/texture.cpp:64-75) -- which is what the CPU oracle
 * computes and every parity test runs.  1: the reference's DEVICE rule, tex2DLod on a texture object with normalised
 * coordinates and the texture's own filter (point / bilinear) and address mode (wrap / clamp / mirror / border)
 * (src/libaten/material/sample_texture.h:18-35, src/libidaten/cuda/cudaTextureResource.cpp:210-236), computed in software
 * after the CUDA programming guide's texture-fetch formulas.  The device rule has no runnable reference on this box:
 * parity unpinned (pinned to an independent restatement of the formulas only). */
ATN_API int atn_set_texture_fetch(AtnRenderer* r, int device_semantics);
/* SceneRenderingConfig::enable_alpha_blending (scene_rendering_config.h:21), also set by the scene blob at upload.  With it
 * on (or for shadow rays that leave a surface of stencil type ALWAYS) a shadow ray whose closest hit is translucent by
 * alpha (or a STENCIL surface) restarts behind that hit, up to 10 lookups -- HitTestToTargetLight,
 * src/libaten/renderer/pathtracing/pathtracing_impl.h:266-351. */
ATN_API int atn_set_alpha_blending(AtnRenderer* r, int enable);
/* Renderer::initSamplerParameter (renderer.h:113-124): per-pixel seeds = aten::getRandom(). */
ATN_API int atn_set_seeds(AtnRenderer* r, const uint32_t* seeds, uint32_t count);

/* ---- frame control ----------------------------------------------------------------------------- */
/* Renderer::GetFrameCount / reset (renderer.h:66-83).  Frame k here == host frame k. */
ATN_API int atn_set_frame(AtnRenderer* r, uint32_t frame);
ATN_API int atn_get_frame(AtnRenderer* r, uint32_t* frame);
/* Renderer::reset + clearing the accumulation image. */
ATN_API int atn_reset(AtnRenderer* r);
/* How a render call folds its samples into the image (the reference's gather, pathtracing_impl.cu:282-324). */
typedef enum AtnGatherMode {
    ATN_GATHER_OVERWRITE = 0,    /* the reference GPU renderer with progressive off (its default, pathtracing.h:103): every
                                    render call overwrites the pixel with the mean of that call's samples, w = 1 */
    ATN_GATHER_PROGRESSIVE = 1,  /* running mean across calls, one step per sample, exactly as FilmProgressive::put
                                    (src/libaten/renderer/film.cpp:61-71): new = (n old + sample) / (n + 1), w = n + 1.
                                    The handle's default; with one sample per call it is also the reference GPU gather
                                    with progressive on */
    ATN_GATHER_SUM = 2           /* no counterpart in the reference: xyz += sample, w += 1 across calls -- the form the
                                    multi-GPU reduce combines (one divide by w after the reduce) */
} AtnGatherMode;
/* PathTracing::SetEnableProgressive (pathtracing.h:127-135): 1 = ATN_GATHER_PROGRESSIVE, 0 = ATN_GATHER_OVERWRITE. */
ATN_API int atn_set_progressive(AtnRenderer* r, int enable);
ATN_API int atn_set_gather_mode(AtnRenderer* r, int mode /* AtnGatherMode */);
/* No counterpart in the reference.  1 (default): a ray whose two nearest primitives lie within the error of the reference's
 * biased box test of each other -- the one case in which the reference's answer depends on its own visit order -- is
 * repeated as the reference's depth-first walk (ThreadedBvhTraverser::Traverse, threaded_bvh_traverser.h:98-304) over the
 * reference's node lists, so that the reported primitive is the reference's.  0: such rays keep the nearest of the tied
 * primitives (a fraction of a percent of the rays of a scene with coincident surfaces; |dt| / t below 1e-4). */
ATN_API int atn_set_exact_ties(AtnRenderer* r, int enable);
/* No counterpart in the reference (its device renderer traces one sample of the frame per pass and sizes its path
 * buffers by the pixel count, renderer.cpp:30-60).  Upper bound on the paths of one wavefront, i.e. on the device memory the
 * path state takes: about 265 bytes per path.  Default 32 Mi paths (8.8 GB): the camera samples of a frame share wavefronts
 * up to that size (all 16 samples of a 1080p frame, 4 of a 4K frame).  A smaller bound trades a few percent of throughput
 * for memory (1080p, 16 spp: 8 Mi paths 2.2 GB -2 %, 4 Mi 1.1 GB -5 %, 2 Mi = one sample per wavefront 0.55 GB -10 %);
 * images and ray counts do not depend on it.  Takes
 * effect at the next atn_upload_scene; at least one sample of the frame always fits. */
ATN_API int atn_set_wavefront_paths(AtnRenderer* r, uint64_t max_paths);
/* Renderer::setHitDistanceLimit (renderer.h:108-111) */
ATN_API int atn_set_hit_distance_limit(AtnRenderer* r, float d);

/* ---- render ------------------------------------------------------------------------------------ */
/* PathTracingImplBase::render(width,height,maxSamples,maxBounce) (pathtracing.cpp:49-89) followed
 * by PathTracing::OnRender (pathtracing.cpp:91-153).  Renders `spp` samples per pixel, sample i of
 * the call using host frame index (frame + i); advances the frame counter by spp.  Asynchronous on
 * the handle's stream: the call returns once the work is queued (it waits only when a bounce tap is installed);
 * atn_readback / atn_synchronize / atn_get_stats wait for it.  The reference synchronises after every kernel
 * (cudautil.h:93-104). */
ATN_API int atn_render(AtnRenderer* r, int32_t width, int32_t height,
                       int32_t spp, int32_t max_bounce, int32_t rr_bounce);
/* Same, restricted to the pixel rectangle [x0,x1) x [y0,y1) and to samples
 * [sample_begin, sample_end) of the frame sequence -- the multi-GPU partition.
 * Does not advance the frame counter. */
ATN_API int atn_render_region(AtnRenderer* r, int32_t width, int32_t height,
                              int32_t x0, int32_t y0, int32_t x1, int32_t y1,
                              int32_t sample_begin, int32_t sample_end,
                              int32_t max_bounce, int32_t rr_bounce);
/* The other multi-GPU partition: the frame is cut into tile x tile squares numbered in scanline order and
 * this call renders the squares g with g % world == rank -- all of them as ONE wavefront, so a rank's share costs what
 * a frame of that many pixels costs -- for samples [sample_begin, sample_end).  Pixels of the other ranks' tiles are
 * not touched.  Does not advance the frame counter; not available with the bounce tap. */
ATN_API int atn_render_tiles(AtnRenderer* r, int32_t width, int32_t height, int32_t tile, int32_t world, int32_t rank,
                             int32_t sample_begin, int32_t sample_end, int32_t max_bounce, int32_t rr_bounce);
/* Row N1 of SURVEY 8f, the first renderer that sits on this path: ambient occlusion.
 * idaten::AORenderer::render (src/libidaten/ao/ao.cpp:9-51, kernels ao.cu:20-121) / aten::AORenderer
 * (src/libaten/renderer/ao/aorenderer.cpp:20-146, ShandeByAO aorenderer_impl.h:33-117): camera ray,
 * closest hit, `num_rays` cosine-distributed rays of length `radius` from the hit, image = AO value in
 * rgb.  The reference hard-codes num_rays = 1, radius = 1.  A primary miss stores 1.0 (GPU kernel
 * semantics, ao.cu:27-33).  Advances the frame counter by spp; accumulates like atn_render. */
ATN_API int atn_render_ao(AtnRenderer* r, int32_t width, int32_t height, int32_t spp, int32_t num_rays, float radius);
/* Row N1 of SURVEY 8f, second renderer on the path: SVGF (spatiotemporal variance-guided filtering).
 * idaten::SVGFPathTracing::render (src/libidaten/svgf/svgf.cpp:28-160; kernels svgf_pt.cu, svgf_tp.cu,
 * svgf_VarianceEstimation.cu, svgf_atrous.cu) / aten::SVGFRenderer::OnRender (src/libaten/renderer/svgf/svgf.cpp:461-639,
 * per-pixel functions svgf_impl.h).  ONE frame per call: one path-traced sample per pixel with the SVGF shade
 * (albedo of the first hit separated, AOVs = geometric normal, clip-space depth, material id), then
 * PrepareForDenoise, temporal reprojection against the previous frame's AOV set (frames > 0), variance
 * estimation, `atrous_iters` a-trous iterations (<= 0: the reference's 5) and the copy into the colour
 * history.  The frame with counter 0 has no history (atn_set_frame(0) / a fresh upload starts a sequence);
 * the counter advances by 1.  The image (atn_readback / atn_device_image) holds the filtered colour with the
 * albedo put back, .w as the reference leaves it (material id x variance).
 * Variance estimation: the reference filters the colour history in place while neighbours read it
 * (svgf_impl.h:438-544), so its result depends on the pixel order; this entry point computes the
 * order-independent form (every pixel reads the unfiltered colours) -- DESIGN.md "SVGF". */
ATN_API int atn_render_svgf(AtnRenderer* r, int32_t width, int32_t height, int32_t max_bounce, int32_t rr_bounce, int32_t atrous_iters);
/* Row N1 of SURVEY 8f, third renderer on the path: ReSTIR (reservoir-based spatiotemporal importance resampling of the
 * direct light at the first hit).  idaten::ReSTIRPathTracing::OnRender (src/libidaten/restir/restir.cpp:47-123, restir.cu,
 * restir_reuse.cu); per-pixel steps = restir_impl.h + aten::ReSTIRRenderer::Shade (src/libaten/renderer/restir/restir.cpp:100-241).
 * ONE frame of one sample per call: camera ray, first hit -> ReSTIRInfo + AOVs (shading normal, clip-space depth, mesh id),
 * one light candidate per pixel, visibility of the kept sample, temporal reuse against the previous frame's reservoirs
 * (frames > 1, through the motion vectors of atn_set_motion_depth), 3x3 spatial reuse, pixel colour; the remaining
 * bounces are the path tracer's.  Frame 0 starts a sequence (atn_set_frame(0)); the counter advances by 1.  The image
 * accumulates like atn_render with progressive mode on, and holds this frame alone with it off. */
ATN_API int atn_render_restir(AtnRenderer* r, int32_t width, int32_t height, int32_t max_bounce, int32_t rr_bounce);
/* The reservoirs the last frame's pixel colours were computed from (AT_NAME::Reservoir, restir_types.h:6-57), five floats
 * per pixel: w_sum, M, y (light index, -1 none), W, target pdf. */
ATN_API int atn_readback_restir(AtnRenderer* r, float* reservoirs, size_t pixel_count);
/* Replaces SVGFPathTracing::SetMotionDepthBuffer / SVGFRenderer::SetMotionDepthBuffer(aten::FBO&) (svgf.cpp:451-459),
 * which read a GL frame buffer: W*H motion vectors in .xy (fractions of the screen, added to the pixel position to
 * find it in the previous frame, svgf_impl.h:325-326).  NULL: static camera (zero motion). */
ATN_API int atn_set_motion_depth(AtnRenderer* r, const AtnVec4* motion_depth, size_t pixel_count);
/* The SVGF AOV set of the last frame rendered (SVGFParams::aovs, svgf_types.h:55-66): normal + depth, albedo +
 * material id, colour history + variance, luminance moments + frames + temporal weight.  Any pointer may be NULL. */
ATN_API int atn_readback_svgf(AtnRenderer* r, AtnVec4* normal_depth, AtnVec4* albedo_meshid, AtnVec4* color_variance,
                              AtnVec4* moment_temporalweight, size_t pixel_count);
/* Blocks until the handle's stream is idle. */
ATN_API int atn_synchronize(AtnRenderer* r);
/* Replaces the CUDA-GL surface the reference writes with surf2Dwrite (pathtracing_impl.cu:282-324):
 * copies the float4 W*H image to host memory. */
ATN_API int atn_readback(AtnRenderer* r, AtnVec4* host_image, size_t pixel_count);
/* The two basic AOV buffers PathTracingImplBase keeps (aov_.normal_depth(), aov_.albedo_meshid(),
 * pathtracing.h:96-98), filled for the first hit of the last sample rendered exactly as the reference's
 * shade / ShadeMiss kernels fill them (FillBasicAOVs with an identity matrix, pathtracing_impl.cu:102-111;
 * FillBasicAOVsIfHitMiss, pathtracing_impl.h:150-158): normal.xyz + depth and albedo.rgb + meshid per
 * pixel.  Either pointer may be NULL. */
ATN_API int atn_readback_aov(AtnRenderer* r, AtnVec4* normal_depth, AtnVec4* albedo_meshid, size_t pixel_count);
/* Device pointer (float4[W*H]) of the image, for NCCL / torch interop; owned by the handle. */
ATN_API int atn_device_image(AtnRenderer* r, void** device_ptr, size_t* pixel_count);
/* Counters of the last call on the handle; waits for that call to finish on the GPU first. */
ATN_API int atn_get_stats(AtnRenderer* r, AtnStats* out);
/* Bit mask, default 0.  1: time the traversal kernels separately with CUDA events (fills AtnStats.trace_ms).  2: count
 * their node fetches and triangle tests (AtnStats.nodes_visited / tris_tested; the counting form of the kernels keeps two
 * more registers per lane alive and is 1.4 % slower, so it is not the default). */
ATN_API int atn_set_profiling(AtnRenderer* r, int enable);
/* CUDA stream of the handle (Renderer::GetCudaStream, renderer.h:113-116) as an opaque pointer. */
ATN_API int atn_get_stream(AtnRenderer* r, void** stream);

/* ---- ray queries (config C5; the traversal the renderer uses, exposed on its own) ------------- */
/* ThreadedBvhTraverser<true>::Traverse<Closest>(isect, ctxt, ray, t_min=1e-9, t_max)
 * (src/libaten/accelerator/threaded_bvh_traverser.h:98-304) for n host rays. */
ATN_API int atn_trace_closest(AtnRenderer* r, const AtnRay* rays, const float* t_max /* may be NULL */,
                              size_t n, AtnHit* hits);
/* Shadow-ray query with the reference semantics of HitTestToTargetLight + scene::hitLight
 * (pathtracing_impl.h:266-351, scene.h:120-133) for a singular light at distance dist[i]:
 * visible[i] = 1 iff nothing is hit closer than dist[i]. */
ATN_API int atn_trace_any(AtnRenderer* r, const AtnRay* rays, const float* dist, size_t n,
                          uint8_t* visible);
/* Device-resident variants (pointers are CUDA device pointers; asynchronous on the stream).  Their CUDA-event time and
 * visit counts are collected by the next atn_get_stats, which waits for the stream. */
ATN_API int atn_trace_closest_device(AtnRenderer* r, const void* rays_dev, const void* t_max_dev,
                                     size_t n, void* hits_dev);
ATN_API int atn_trace_any_device(AtnRenderer* r, const void* rays_dev, const void* dist_dev,
                                 size_t n, void* visible_dev);

/* No counterpart in the reference.  Same query as atn_trace_closest (host rays in, host hits out,
 * t_max may be NULL) for callers that must not allocate device memory: hosts that run inside a fixed
 * device-memory budget which is already spent (so a cudaMalloc may fail), and frame loops that cannot
 * take cudaMalloc's device-wide synchronisation.  Rays and hits are staged in an array that is part of
 * the library image on every GPU it is loaded on (~4.2 MB, held until the process exits) instead of
 * buffers grown on demand; the call performs no cudaMalloc / cudaFree and calls no allocating routine.
 *   - at most 65536 rays per call; more is refused with ATN_ERR_TOO_LARGE (split the batch);
 *   - the first call on each GPU looks the array's device address up (cudaGetSymbolAddress); make that
 *     first call before any stream capture begins;
 *   - all handles and threads on one GPU share the one array: calls that overlap on the GPU race and
 *     corrupt each other's rays and hits -- serialise them (one call at a time per GPU).
 * Synchronous: returns when `hits` is filled. */
ATN_API int atn_trace_closest_static(AtnRenderer* r, const AtnRay* rays, const float* t_max /* may be NULL */,
                                     size_t n, AtnHit* hits);

/* No counterpart in the reference: a measurement aid for the roofline figures (bench.py).  Read bandwidth in GB/s of a
 * device buffer of `bytes` walked `passes` times by ONE persistent launch with 256-bit loads: with bytes below the L2 size
 * this is the L2 read bandwidth, far above it the HBM read bandwidth.  Allocates and frees its own buffer. */
ATN_API int atn_probe_read_bandwidth(int device_ordinal, size_t bytes, int passes, float* gb_per_s, float* ms /* may be NULL */);

/* ---- debug / parity taps ------------------------------------------------------------------------ */
/* Per-bounce dump of the wavefront state for parity tests: after atn_render with a tap installed,
 * `records` holds one AtnBounceRecord per (pixel, bounce) for the LAST sample rendered.
 * Host memory; bounce_capacity records per pixel. Pass NULL to remove. */
typedef struct AtnBounceRecord {
    AtnRay ray;            /* ray traced at this bounce */
    AtnHit hit;            /* closest hit found */
    float shadow_org[3]; float shadow_dist;
    float shadow_dir[3]; int32_t shadow_active;   /* 1 if a shadow ray was cast */
    float light_contrib[3]; int32_t shadow_visible;
    float throughput[3]; float pdfb;              /* after the bounce */
    float contrib[3]; int32_t flags;              /* bit0 valid, bit1 terminated after bounce */
} AtnBounceRecord;
ATN_API int atn_set_bounce_tap(AtnRenderer* r, AtnBounceRecord* records, int32_t bounce_capacity);

#ifdef __cplusplus
}
#endif
#endif /* ATEN_B200_H_ */
