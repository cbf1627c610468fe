// TEST INFRASTRUCTURE (oracle/_ref): headless driver around the UNMODIFIED reference CPU path tracer.
//
// Links against the reference's own libaten / libatenscene sources, compiled where they lie by
// oracle/build_ref.py.  It does what src/host_renderer/main.cpp:42-229 does, without a window:
// build a scene through the reference's loaders and SBVH builder, render with
// aten::PathTracing (src/libaten/renderer/pathtracing/pathtracing.cpp:269-366), and write results
// as raw little-endian files that tests/ and bench.py read.  It also
//   * dumps the flattened scene through the drop-in shim (include/idaten_b200_shim.h) -- the same
//     arrays a maintainer would hand to libaten_b200.so;
//   * replays aten::PathTracing::radiance() bounce by bounce using the reference's PUBLIC functions
//     (GeneratePath, BvhTraverser::Traverse, PathTracing::shade, HitShadowRay, ShadeMiss) so that
//     every ray / hit / shadow ray / contribution can be compared per bounce; the replay is checked
//     bit-exact against the stock render() on every run of `trace`;
//   * answers closest-hit queries for arbitrary rays.
// Nothing here is product code; only tests/, __graft_entry__.smoke() and bench.py's reference arm
// execute it.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <unistd.h>
#include <dirent.h>
#include <omp.h>

#include "aten.h"
#include "atenscene.h"
#include "renderer/pathtracing/pathtracing_impl.h"
#include "renderer/svgf/svgf.h"
#include "renderer/restir/restir.h"
#include "renderer/restir/restir_impl.h"

#include "../include/idaten_b200_shim.h"

namespace {

std::string g_asset;   // directory that holds cornellbox/, sponza/, ...

std::shared_ptr<aten::material> MakeMtrl(
    std::string_view name, aten::context& ctxt, aten::MaterialType type, const aten::vec3& clr,
    aten::texture* albedo = nullptr, aten::texture* nml = nullptr)
{
    aten::MaterialParameter p;
    p.type = type;
    p.baseColor = clr;
    return ctxt.CreateMaterialWithMaterialParameter(name, p, albedo, nml, nullptr);
}

std::shared_ptr<aten::material> MakeMtrlP(
    std::string_view name, aten::context& ctxt, aten::MaterialType type, aten::MaterialParameter p)
{
    p.type = type;
    return ctxt.CreateMaterialWithMaterialParameter(name, p, nullptr, nullptr, nullptr);
}

std::string FindNoCase(const std::string& dir, const std::string& file)
{
    auto lower = [](std::string v) { for (auto& ch : v) ch = (char)tolower(ch); return v; };
    DIR* d = opendir(dir.c_str());
    std::string found = dir + "/" + file;
    if (!d) return found;
    const std::string want = lower(file);
    while (auto* e = readdir(d)) {
        if (lower(e->d_name) == want) { found = dir + "/" + e->d_name; break; }
    }
    closedir(d);
    return found;
}

// Analytic sky written into an aten::texture: the snapshot has no asset/envmap/*.hdr.
std::shared_ptr<aten::texture> MakeSky(aten::context& ctxt, int32_t w, int32_t h)
{
    auto tex = ctxt.CreateTexture(w, h, 4, "synthetic_sky");
    for (int32_t y = 0; y < h; y++) {
        for (int32_t x = 0; x < w; x++) {
            const float v = (y + 0.5f) / h;          // 0 = bottom (-y) ... 1 = top (+y)
            const float u = (x + 0.5f) / w;
            const float up = v;                       // horizon at 0.5
            aten::vec3 c = up > 0.5f
                ? aten::vec3(0.35f, 0.55f, 0.95f) * (0.6f + 0.8f * (up - 0.5f)) + aten::vec3(0.9f) * (1.0f - up) * 0.8f
                : aten::vec3(0.25f, 0.22f, 0.2f) * (0.3f + up);
            // a soft sun around u = 0.2, v = 0.8
            const float du = u - 0.2f, dv = v - 0.8f;
            const float s = std::exp(-(du * du + dv * dv) * 400.0f);
            c += aten::vec3(12.0f, 10.5f, 8.0f) * s;
            tex->PutByXYcoord(x, y, aten::vec4(c, 1.0f));
        }
    }
    return tex;
}

struct Harness {
    aten::context ctxt;
    aten::AcceleratedScene<aten::sbvh> scene;
    aten::PinholeCamera camera;
    std::shared_ptr<aten::texture> envmap;
    int32_t W{ 0 }, H{ 0 };

    void SetCamera(const aten::vec3& from, const aten::vec3& at, float fov)
    {
        camera.Initalize(from, at, aten::vec3(0, 1, 0), fov, 0.1f, 10000.0f, W, H);
    }

    // ---- scenes -------------------------------------------------------------------------------
    // C1: the reference's ObjCornellBoxScene (src/common/scenedefs.cpp:732-803), restated through
    // the same public calls so that asset paths are not tied to the cwd.
    void BuildCornell(float light_dy = 0.0f)
    {
        auto emit = MakeMtrl("light", ctxt, aten::MaterialType::Emissive, aten::vec3(1.0f));
        auto objs = aten::ObjLoader::Load(
            g_asset + "/cornellbox/orig.obj", ctxt,
            [&](std::string_view name, aten::context& c, aten::MaterialType type, const aten::vec3& clr,
                const std::string&, const std::string&) -> std::shared_ptr<aten::material> {
                if (name == "shortBox" || name == "floor") {
                    aten::MaterialParameter p;
                    p.baseColor = aten::vec3(0.7f, 0.6f, 0.5f);
                    p.standard.roughness = 0.1f;
                    p.standard.ior = 0.01f;
                    return MakeMtrlP(name, c, name == "floor" ? aten::MaterialType::GGX : aten::MaterialType::Specular, p);
                }
                return MakeMtrl(name, c, type, clr);
            },
            nullptr, true, true);
        auto light = aten::TransformableFactory::createInstance<aten::PolygonObject>(
            ctxt, objs[0], aten::vec3(0.0f, light_dy, 0.0f), aten::vec3(0.0f), aten::vec3(1.0f));
        scene.add(light);
        ctxt.AddLight(std::make_shared<aten::AreaLight>(light, emit->param().baseColor, 200.0f));
        for (size_t i = 1; i < objs.size(); i++) {
            scene.add(aten::TransformableFactory::createInstance<aten::PolygonObject>(ctxt, objs[i], aten::mat4::Identity));
        }
        SetCamera(aten::vec3(0.f, 1.f, 3.f), aten::vec3(0.f, 1.f, 0.f), 45.0f);
        ctxt.scene_rendering_config.bg = AT_NAME::Background::CreateBackgroundResource(nullptr, aten::vec4(0.0f));
    }

    // Cornell with other materials on the boxes / walls: exercises Beckmann, Disney, refraction and
    // point / spot / directional lights on the same 32-triangle geometry.
    void BuildCornellMaterials()
    {
        auto emit = MakeMtrl("light", ctxt, aten::MaterialType::Emissive, aten::vec3(1.0f));
        auto objs = aten::ObjLoader::Load(
            g_asset + "/cornellbox/orig.obj", ctxt,
            [&](std::string_view name, aten::context& c, aten::MaterialType type, const aten::vec3& clr,
                const std::string&, const std::string&) -> std::shared_ptr<aten::material> {
                aten::MaterialParameter p;
                if (name == "shortBox") {
                    p.baseColor = aten::vec3(0.95f, 0.95f, 0.95f);
                    p.standard.ior = 1.5f;
                    return MakeMtrlP(name, c, aten::MaterialType::Refraction, p);
                }
                if (name == "tallBox") {
                    p.baseColor = aten::vec3(0.82f, 0.67f, 0.16f);
                    p.standard.roughness = 0.3f; p.standard.metallic = 0.6f; p.standard.specular = 0.7f;
                    p.standard.subsurface = 0.1f; p.standard.sheen = 0.3f; p.standard.clearcoat = 0.8f;
                    p.standard.clearcoatGloss = 0.7f; p.standard.specularTint = 0.2f; p.standard.sheenTint = 0.4f;
                    return MakeMtrlP(name, c, aten::MaterialType::Disney, p);
                }
                if (name == "floor") {
                    p.baseColor = aten::vec3(0.7f, 0.6f, 0.5f);
                    p.standard.roughness = 0.25f; p.standard.ior = 1.8f;
                    return MakeMtrlP(name, c, aten::MaterialType::Beckman, p);
                }
                if (name == "backWall") {
                    p.baseColor = aten::vec3(0.6f, 0.7f, 0.8f);
                    p.standard.roughness = 0.4f; p.standard.ior = 2.5f;
                    return MakeMtrlP(name, c, aten::MaterialType::GGX, p);
                }
                return MakeMtrl(name, c, type, clr);
            },
            nullptr, true, true);
        auto light = aten::TransformableFactory::createInstance<aten::PolygonObject>(
            ctxt, objs[0], aten::vec3(0.0f), aten::vec3(0.0f), aten::vec3(1.0f));
        scene.add(light);
        ctxt.AddLight(std::make_shared<aten::AreaLight>(light, emit->param().baseColor, 120.0f));
        ctxt.AddLight(std::make_shared<aten::PointLight>(aten::vec3(-0.6f, 1.6f, 0.4f), aten::vec3(1.0f, 0.8f, 0.6f), 6.0f));
        ctxt.AddLight(std::make_shared<aten::SpotLight>(
            aten::vec3(0.6f, 1.8f, 0.6f), aten::vec3(-0.3f, -1.0f, -0.4f), aten::vec3(0.6f, 0.8f, 1.0f), 12.0f,
            aten::Deg2Rad(30.0f), aten::Deg2Rad(50.0f)));
        ctxt.AddLight(std::make_shared<aten::DirectionalLight>(aten::vec3(-0.2f, -1.0f, -0.6f), aten::vec3(1.0f), 0.3f));
        for (size_t i = 1; i < objs.size(); i++) {
            scene.add(aten::TransformableFactory::createInstance<aten::PolygonObject>(ctxt, objs[i], aten::mat4::Identity));
        }
        SetCamera(aten::vec3(0.f, 1.f, 3.f), aten::vec3(0.f, 1.f, 0.f), 45.0f);
        ctxt.scene_rendering_config.bg = AT_NAME::Background::CreateBackgroundResource(nullptr, aten::vec4(0.05f, 0.06f, 0.08f, 1.0f));
    }

    // Cornell box lit by a SPHERE area light (sphere::SamplePosAndNormal + sphere::hit, src/libaten/geometry/sphere.cpp:30-150,
    // AreaLight::sample's "geometric form" branch, src/libaten/light/arealight.h:88-130) next to the polygon light, with a
    // second, non-emissive sphere object in the top layer: the reference's closest-hit traversal never reports a sphere
    // (a top-layer leaf without an external tree holds nothing traversable, threaded_bvh_traverser.h:142-220), so that
    // sphere must stay invisible and must not occlude shadow rays either.
    void BuildCornellSphere()
    {
        auto emit = MakeMtrl("light", ctxt, aten::MaterialType::Emissive, aten::vec3(1.0f));
        auto objs = aten::ObjLoader::Load(
            g_asset + "/cornellbox/orig.obj", ctxt,
            [&](std::string_view name, aten::context& c, aten::MaterialType type, const aten::vec3& clr,
                const std::string&, const std::string&) -> std::shared_ptr<aten::material> {
                return MakeMtrl(name, c, type, clr);
            },
            nullptr, true, true);
        auto light = aten::TransformableFactory::createInstance<aten::PolygonObject>(
            ctxt, objs[0], aten::vec3(0.0f), aten::vec3(0.0f), aten::vec3(1.0f));
        scene.add(light);
        ctxt.AddLight(std::make_shared<aten::AreaLight>(light, emit->param().baseColor, 60.0f));
        auto emit2 = MakeMtrl("sphere_light", ctxt, aten::MaterialType::Emissive, aten::vec3(1.0f, 0.8f, 0.5f));
        auto ball = aten::TransformableFactory::createSphere(ctxt, aten::vec3(-0.45f, 1.45f, 0.35f), 0.18f, emit2);
        scene.add(ball);
        ctxt.AddLight(std::make_shared<aten::AreaLight>(ball, emit2->param().baseColor, 150.0f));
        auto grey = MakeMtrl("sphere_grey", ctxt, aten::MaterialType::Diffuse, aten::vec3(0.5f));
        scene.add(aten::TransformableFactory::createSphere(ctxt, aten::vec3(0.35f, 0.9f, 0.4f), 0.3f, grey));
        for (size_t i = 1; i < objs.size(); i++) {
            scene.add(aten::TransformableFactory::createInstance<aten::PolygonObject>(ctxt, objs[i], aten::mat4::Identity));
        }
        SetCamera(aten::vec3(0.f, 1.f, 3.f), aten::vec3(0.f, 1.f, 0.f), 45.0f);
        ctxt.scene_rendering_config.bg = AT_NAME::Background::CreateBackgroundResource(nullptr, aten::vec4(0.0f));
    }

    // The reference's AlphaBlendedObjCornellBoxScene (src/common/scenedefs.cpp:888-926): Cornell box whose back wall, left
    // wall and tall box are translucent by alpha, with SceneRenderingConfig::enable_alpha_blending on -- shadow rays pass
    // through those surfaces in up to 10 lookups (HitTestToTargetLight, pathtracing_impl.h:266-351).  The floor gets
    // stencil type ALWAYS and the short box STENCIL, so that the stencil branch of the same loop is exercised too.
    void BuildCornellAlpha()
    {
        ctxt.scene_rendering_config.enable_alpha_blending = true;
        auto emit = MakeMtrl("light", ctxt, aten::MaterialType::Emissive, aten::vec3(1.0f));
        auto objs = aten::ObjLoader::Load(
            g_asset + "/cornellbox/orig.obj", ctxt,
            [&](std::string_view name, aten::context& c, aten::MaterialType type, const aten::vec3& clr,
                const std::string&, const std::string&) -> std::shared_ptr<aten::material> {
                auto m = MakeMtrl(name, c, type, clr);
                if (name == "backWall") m->param().baseColor.a = 0.0f;
                if (name == "leftWall") m->param().baseColor.a = 0.5f;
                if (name == "tallBox") m->param().baseColor.a = 0.25f;
                if (name == "floor") m->param().stencil_type = aten::StencilType::ALWAYS;
                if (name == "shortBox") m->param().stencil_type = aten::StencilType::STENCIL;
                return m;
            },
            nullptr, true, true);
        auto light = aten::TransformableFactory::createInstance<aten::PolygonObject>(
            ctxt, objs[0], aten::vec3(0.0f), aten::vec3(0.0f), aten::vec3(1.0f));
        scene.add(light);
        ctxt.AddLight(std::make_shared<aten::AreaLight>(light, emit->param().baseColor, 100.0f));
        // lights behind the translucent walls: their shadow rays have to pass the walls
        ctxt.AddLight(std::make_shared<aten::PointLight>(aten::vec3(-2.5f, 1.2f, 0.3f), aten::vec3(1.0f, 0.8f, 0.6f), 40.0f));
        ctxt.AddLight(std::make_shared<aten::DirectionalLight>(aten::vec3(0.1f, -0.3f, 1.0f), aten::vec3(0.6f, 0.8f, 1.0f), 1.5f));
        for (size_t i = 1; i < objs.size(); i++) {
            scene.add(aten::TransformableFactory::createInstance<aten::PolygonObject>(ctxt, objs[i], aten::mat4::Identity));
        }
        SetCamera(aten::vec3(0.f, 1.f, 3.f), aten::vec3(0.f, 1.f, 0.f), 45.0f);
        ctxt.scene_rendering_config.bg = AT_NAME::Background::CreateBackgroundResource(nullptr, aten::vec4(0.0f));
    }

    // C2 stand-in: asset/sponza/sponza_lod.obj + the shipped sponza_lod.sbvh (the full-resolution
    // Sponza OBJ/SBVH are stripped from the snapshot).  Materials from the .mtl
    // (Lambert), arches/columns switched to GGX; lit by the synthetic-sky IBL plus one area light.
    // `textured` binds the .mtl's albedo / normal maps (C3 stand-in).
    void BuildSponza(bool textured, bool atrium_camera)
    {
        auto objs = aten::ObjLoader::Load(
            g_asset + "/sponza/sponza_lod.obj", ctxt,
            [&](std::string_view name, aten::context& c, aten::MaterialType type, const aten::vec3& clr,
                const std::string& albedo, const std::string& nml) -> std::shared_ptr<aten::material> {
                std::shared_ptr<aten::texture> at, nt;
                if (textured) {
                    // sponza.mtl names its maps in another letter case than the files on disk
                    if (!albedo.empty()) at = aten::ImageLoader::load(FindNoCase(g_asset + "/sponza", albedo), c);
                    if (!nml.empty()) nt = aten::ImageLoader::load(FindNoCase(g_asset + "/sponza", nml), c);
                }
                const std::string n(name);
                if (n.find("stup") != std::string::npos || n.find("luk") != std::string::npos) {
                    aten::MaterialParameter p;
                    p.baseColor = clr;
                    p.standard.roughness = 0.35f;
                    p.standard.ior = 1.6f;
                    p.type = aten::MaterialType::GGX;
                    return c.CreateMaterialWithMaterialParameter(name, p, at.get(), nt.get(), nullptr);
                }
                return MakeMtrl(name, c, type, clr, at.get(), nt.get());
            });
        objs[0]->importInternalAccelTree(g_asset + "/sponza/sponza_lod.sbvh", ctxt, 0);
        auto sponza = aten::TransformableFactory::createInstance<aten::PolygonObject>(ctxt, objs[0], aten::mat4::Identity);
        scene.add(sponza);

        // One emissive quad (asset/cornellbox light geometry is tiny; build from the cornell light obj).
        auto emit = MakeMtrl("light", ctxt, aten::MaterialType::Emissive, aten::vec3(1.0f, 0.95f, 0.85f));
        auto lobjs = aten::ObjLoader::Load(
            g_asset + "/cornellbox/orig.obj", ctxt,
            [&](std::string_view name, aten::context& c, aten::MaterialType type, const aten::vec3& clr,
                const std::string&, const std::string&) -> std::shared_ptr<aten::material> {
                return MakeMtrl(name, c, type, clr);
            },
            nullptr, true, true);
        auto light = aten::TransformableFactory::createInstance<aten::PolygonObject>(
            ctxt, lobjs[0], aten::vec3(0.0f, 10.0f, 0.0f), aten::vec3(0.0f), aten::vec3(1.0f));
        scene.add(light);
        ctxt.AddLight(std::make_shared<aten::AreaLight>(light, emit->param().baseColor, 4000.0f));

        if (atrium_camera) SetCamera(aten::vec3(-12.0f, 4.0f, 0.5f), aten::vec3(0.0f, 5.0f, 0.0f), 60.0f);
        else SetCamera(aten::vec3(0.f, 1.f, 3.f), aten::vec3(0.f, 1.f, 0.f), 45.0f);

        envmap = MakeSky(ctxt, 512, 256);
        ctxt.scene_rendering_config.bg = AT_NAME::Background::CreateBackgroundResource(envmap);
    }

    // C4 stand-in: transformed instances of teapot / suzanne (+ dragon when `with_dragon`), Disney
    // and refraction materials (parameters follow DisneyBrdfScene, src/common/scenedefs.cpp:1506-1545).
    void BuildInstanced(bool with_dragon)
    {
        aten::MaterialParameter disney;
        disney.baseColor = aten::vec3(0.82f, 0.67f, 0.16f);
        disney.standard.roughness = 0.3f; disney.standard.specular = 0.5f; disney.standard.metallic = 0.5f;
        disney.standard.subsurface = 0.2f; disney.standard.sheen = 0.2f; disney.standard.clearcoat = 0.6f;
        auto m_disney = MakeMtrlP("disney", ctxt, aten::MaterialType::Disney, disney);
        aten::MaterialParameter glass;
        glass.baseColor = aten::vec3(0.9f, 0.95f, 1.0f);
        glass.standard.ior = 1.33f;
        auto m_glass = MakeMtrlP("glass", ctxt, aten::MaterialType::Refraction, glass);
        aten::MaterialParameter ggx;
        ggx.baseColor = aten::vec3(0.7f, 0.3f, 0.3f); ggx.standard.roughness = 0.2f; ggx.standard.ior = 1.9f;
        auto m_ggx = MakeMtrlP("ggx", ctxt, aten::MaterialType::GGX, ggx);
        auto m_floor = MakeMtrl("floor", ctxt, aten::MaterialType::Diffuse, aten::vec3(0.6f, 0.6f, 0.6f));

        std::shared_ptr<aten::material> cur;
        auto cb = [&](std::string_view, aten::context&, aten::MaterialType, const aten::vec3&,
                      const std::string&, const std::string&) -> std::shared_ptr<aten::material> { return cur; };
        auto find = [&](std::string_view, const aten::context&) -> std::shared_ptr<aten::material> { return cur; };

        cur = m_disney;
        auto teapot = aten::ObjLoader::Load(g_asset + "/teapot/teapot.obj", ctxt, cb, find, false, true);
        cur = m_glass;
        auto suzanne = aten::ObjLoader::Load(g_asset + "/suzanne/suzanne.obj", ctxt, cb, find, false, true);
        std::vector<std::shared_ptr<aten::PolygonObject>> dragon;
        if (with_dragon) {
            cur = m_ggx;
            dragon = aten::ObjLoader::Load(g_asset + "/dragon/dragon.obj", ctxt, cb, find, false, true);
        }
        cur = m_floor;
        auto plane = aten::ObjLoader::Load(g_asset + "/plane/plane.obj", ctxt, cb, find, false, true);

        // NOTE: instance rotation is RotX(rx) RotX(ry) RotX(rz); scale is baked into vertices.
        scene.add(aten::TransformableFactory::createInstance<aten::PolygonObject>(
            ctxt, plane[0], aten::vec3(0.0f, 0.0f, 0.0f), aten::vec3(-AT_MATH_PI_HALF, 0.0f, 0.0f), aten::vec3(1.0f)));
        for (int i = 0; i < 3; i++) {
            scene.add(aten::TransformableFactory::createInstance<aten::PolygonObject>(
                ctxt, teapot[0], aten::vec3(-2.5f + 2.5f * i, 0.0f, -1.0f + 0.3f * i),
                aten::vec3(0.2f * i, 0.0f, 0.0f), aten::vec3(1.0f)));
        }
        for (int i = 0; i < 2; i++) {
            scene.add(aten::TransformableFactory::createInstance<aten::PolygonObject>(
                ctxt, suzanne[0], aten::vec3(-1.5f + 3.0f * i, 1.2f, 1.2f),
                aten::vec3(-0.3f + 0.5f * i, 0.0f, 0.0f), aten::vec3(1.0f)));
        }
        if (with_dragon) {
            scene.add(aten::TransformableFactory::createInstance<aten::PolygonObject>(
                ctxt, dragon[0], aten::vec3(0.0f, 0.5f, 0.0f), aten::vec3(0.0f), aten::vec3(1.0f)));
        }
        ctxt.AddLight(std::make_shared<aten::PointLight>(aten::vec3(3.0f, 5.0f, 4.0f), aten::vec3(1.0f, 0.9f, 0.8f), 400.0f));
        ctxt.AddLight(std::make_shared<aten::SpotLight>(
            aten::vec3(-3.0f, 6.0f, 3.0f), aten::vec3(0.5f, -1.0f, -0.5f), aten::vec3(0.7f, 0.8f, 1.0f), 600.0f,
            aten::Deg2Rad(25.0f), aten::Deg2Rad(45.0f)));
        SetCamera(aten::vec3(0.0f, 3.0f, 8.0f), aten::vec3(0.0f, 1.0f, 0.0f), 45.0f);
        envmap = MakeSky(ctxt, 512, 256);
        ctxt.scene_rendering_config.bg = AT_NAME::Background::CreateBackgroundResource(envmap);
    }

    // C5 stand-in: a k x k grid of transformed instances of the 100 000-triangle dragon over a ground
    // plane (k = 10: 10 M instanced triangles through the two-level structure), one SBVH built once by
    // the reference's builder.  A 10 M-triangle single mesh would take the reference builder hours.
    void BuildDragonGrid(int k)
    {
        aten::MaterialParameter ggx;
        ggx.baseColor = aten::vec3(0.75f, 0.65f, 0.4f); ggx.standard.roughness = 0.3f; ggx.standard.ior = 1.8f;
        auto m_ggx = MakeMtrlP("ggx", ctxt, aten::MaterialType::GGX, ggx);
        auto m_floor = MakeMtrl("floor", ctxt, aten::MaterialType::Diffuse, aten::vec3(0.6f, 0.6f, 0.6f));
        std::shared_ptr<aten::material> cur;
        auto cb = [&](std::string_view, aten::context&, aten::MaterialType, const aten::vec3&,
                      const std::string&, const std::string&) -> std::shared_ptr<aten::material> { return cur; };
        auto find = [&](std::string_view, const aten::context&) -> std::shared_ptr<aten::material> { return cur; };
        cur = m_ggx;
        auto dragon = aten::ObjLoader::Load(g_asset + "/dragon/dragon.obj", ctxt, cb, find, false, true);
        cur = m_floor;
        auto plane = aten::ObjLoader::Load(g_asset + "/plane/plane.obj", ctxt, cb, find, false, true);
        const auto bb = dragon[0]->getBoundingbox();
        const auto size = bb.size();
        const float pitch = 1.3f * std::max(size.x, size.z);
        for (int j = 0; j < k; j++) {
            for (int i = 0; i < k; i++) {
                scene.add(aten::TransformableFactory::createInstance<aten::PolygonObject>(
                    ctxt, dragon[0],
                    aten::vec3((i - 0.5f * (k - 1)) * pitch, -bb.minPos().y, (j - 0.5f * (k - 1)) * pitch),
                    aten::vec3(0.03f * ((i * 7 + j * 3) % 11), 0.0f, 0.0f), aten::vec3(1.0f)));
            }
        }
        scene.add(aten::TransformableFactory::createInstance<aten::PolygonObject>(
            ctxt, plane[0], aten::vec3(0.0f), aten::vec3(-AT_MATH_PI_HALF, 0.0f, 0.0f), aten::vec3(1.0f)));
        const float ext = 0.5f * k * pitch;
        ctxt.AddLight(std::make_shared<aten::PointLight>(aten::vec3(0.0f, 3.0f * ext, 0.5f * ext), aten::vec3(1.0f, 0.95f, 0.9f), 40.0f * ext * ext));
        SetCamera(aten::vec3(0.0f, 1.2f * ext, 1.8f * ext), aten::vec3(0.0f, 0.0f, 0.0f), 45.0f);
        envmap = MakeSky(ctxt, 512, 256);
        ctxt.scene_rendering_config.bg = AT_NAME::Background::CreateBackgroundResource(envmap);
    }

    bool Build(const std::string& name, int32_t w, int32_t h)
    {
        W = w; H = h;
        aten::initSampler(W, H);   // std::mt19937(0), src/libaten/sampler/sampler.cpp:8-18
        if (name == "cornell") BuildCornell();
        else if (name == "cornell_mtrl") BuildCornellMaterials();
        else if (name == "cornell_sphere") BuildCornellSphere();
        else if (name == "cornell_alpha") BuildCornellAlpha();
        else if (name == "sponza") BuildSponza(false, false);
        else if (name == "sponza_atrium") BuildSponza(false, true);
        else if (name == "sponza_tex") BuildSponza(true, true);
        else if (name == "instanced") BuildInstanced(false);
        else if (name == "instanced_dragon") BuildInstanced(true);
        else if (name == "cornell_moved") BuildCornell(-0.3f);
        else if (name == "dragon_grid4") BuildDragonGrid(4);
        else if (name == "dragon_grid10") BuildDragonGrid(10);
        else { fprintf(stderr, "unknown scene %s\n", name.c_str()); return false; }

        scene.build(ctxt);
        if (envmap) {
            auto ibl = std::make_shared<aten::ImageBasedLight>(ctxt.scene_rendering_config.bg, ctxt);
            scene.addImageBasedLight(ctxt, ibl);
            // device_renderer sets avgIllum from the IBL (src/device_renderer/main.cpp:123-126),
            // host_renderer leaves 1.0.  The harness fixes it to the IBL value on both sides.
            ctxt.scene_rendering_config.bg.avgIllum = ibl->getAvgIlluminace();
        }
        ctxt.CopyBvhNodes(scene.getAccel()->getNodes());
        ctxt.UpdateSceneBoundingBox(scene);
        camera.update();
        return true;
    }

    aten::Destination MakeDst(aten::Film* film, int32_t depth, int32_t rr)
    {
        aten::Destination dst;
        dst.width = W; dst.height = H;
        dst.maxDepth = depth; dst.russianRouletteDepth = rr; dst.sample = 1;
        dst.buffer = film;
        return dst;
    }
};

void WriteFile(const std::string& path, const void* p, size_t n)
{
    FILE* fp = fopen(path.c_str(), "wb");
    if (!fp) { fprintf(stderr, "cannot write %s\n", path.c_str()); exit(2); }
    fwrite(p, 1, n, fp);
    fclose(fp);
}

// Exposes the protected frame counter so that the replay can follow stock frames.
struct CountingPathTracing : public aten::PathTracing {
    void SetFrame(uint32_t f) { frame_count_ = f; }
};

struct Counters {
    uint64_t closest{ 0 }, shadow{ 0 };
};

// Bounce-by-bounce replay of aten::PathTracing::radiance() (pathtracing.cpp:23-89) for one frame,
// recording AtnBounceRecord per (pixel, bounce).
void Replay(Harness& hs, uint32_t frame, int32_t max_depth, int32_t rr_depth,
            std::vector<AtnBounceRecord>* records, std::vector<aten::vec4>* image, Counters* cnt)
{
    const int32_t W = hs.W, H = hs.H;
    if (rr_depth > max_depth) rr_depth = max_depth - 1;    // pathtracing.cpp:281-283
    aten::PathHost ph;
    ph.init(W, H);
    ph.Clear(frame);
    std::vector<aten::ray> rays(W * H);
    std::vector<aten::ShadowRay> shadow(W * H);
    if (records) records->assign((size_t)W * H * max_depth, AtnBounceRecord{});
    if (image) image->assign((size_t)W * H, aten::vec4(0.0f));
    const auto& cam = hs.camera.param();
    uint64_t n_closest = 0, n_shadow = 0;
#pragma omp parallel for reduction(+:n_closest, n_shadow)
    for (int32_t y = 0; y < H; y++) {
        for (int32_t x = 0; x < W; x++) {
            const int32_t idx = y * W + x;
            const auto rnd = aten::getRandom(idx);
            AT_NAME::GeneratePath(rays[idx], idx, x, y, W, H, 0, frame, ph.paths, cam, rnd);
            ph.paths.contrib[idx].contrib = aten::vec3(0);
            int32_t depth = 0;
            while (depth < max_depth) {
                aten::Intersection isect;
                const aten::ray ray = rays[idx];
                ph.paths.attrib[idx].attr.isHit = false;
                bool is_hit = aten::BvhTraverser::Traverse<aten::IntersectType::Closest>(
                    isect, hs.ctxt, ray, AT_MATH_EPSILON, AT_MATH_INF);
                n_closest++;
                AtnBounceRecord rec{};
                memcpy(&rec.ray, &ray, sizeof(ray));
                memcpy(&rec.hit, &isect, sizeof(isect));
                rec.flags = 1;
                bool will_continue = true;
                if (is_hit) {
                    ph.paths.attrib[idx].attr.isHit = true;
                    aten::PathTracing::shade(idx, ph.paths, hs.ctxt, rays.data(), shadow.data(), isect, &hs.scene, rr_depth, depth);
                    const auto& mtrl = hs.ctxt.GetMaterial(isect.mtrlid);
                    const auto& sr = shadow[idx];
                    const bool traced = sr.isActive && !ph.paths.attrib[idx].attr.is_terminated;
                    if (traced) n_shadow++;
                    bool vis = AT_NAME::HitShadowRay(depth, hs.ctxt, mtrl, ph.paths.attrib[idx], ph.paths.contrib[idx], sr);
                    rec.shadow_active = traced ? 1 : 0;
                    if (traced) {
                        rec.shadow_org[0] = sr.rayorg.x; rec.shadow_org[1] = sr.rayorg.y; rec.shadow_org[2] = sr.rayorg.z;
                        rec.shadow_dist = sr.distToLight;
                        rec.shadow_dir[0] = sr.raydir.x; rec.shadow_dir[1] = sr.raydir.y; rec.shadow_dir[2] = sr.raydir.z;
                        rec.light_contrib[0] = sr.lightcontrib.x; rec.light_contrib[1] = sr.lightcontrib.y; rec.light_contrib[2] = sr.lightcontrib.z;
                    }
                    rec.shadow_visible = vis ? 1 : 0;
                    will_continue = !ph.paths.attrib[idx].attr.is_terminated;
                }
                else {
                    AT_NAME::ShadeMiss(idx, x, y, W, H, depth, hs.ctxt, cam, ph.paths, rays[idx]);
                    will_continue = false;
                }
                const auto& tp = ph.paths.throughput[idx];
                rec.throughput[0] = tp.throughput.x; rec.throughput[1] = tp.throughput.y; rec.throughput[2] = tp.throughput.z;
                rec.pdfb = tp.pdfb;
                const auto& c = ph.paths.contrib[idx].contrib;
                rec.contrib[0] = c.x; rec.contrib[1] = c.y; rec.contrib[2] = c.z;
                if (!will_continue) rec.flags |= 2;
                if (records) (*records)[(size_t)idx * max_depth + depth] = rec;
                if (!will_continue) break;
                depth++;
            }
            if (image) {
                const auto& c = ph.paths.contrib[idx].contrib;
                (*image)[idx] = aten::vec4(c.x, c.y, c.z, 1.0f);
            }
        }
    }
    if (cnt) { cnt->closest += n_closest; cnt->shadow += n_shadow; }
}

// aten::SVGFRenderer takes its motion / depth buffer from a GL frame buffer object
// (SVGFRenderer::SetMotionDepthBuffer(aten::FBO&), svgf.cpp:451-459).  There is no GL here, so the
// harness writes the vector the renderer reads -- SVGFParams::motion_depth_buffer inside the private
// member params_ -- through an explicit-instantiation accessor (access checks do not apply to the
// arguments of an explicit template instantiation).
using SvgfParamsT = AT_NAME::SVGFParams<std::vector<aten::vec4>>;
struct SvgfParamsTag { using type = SvgfParamsT aten::SVGFRenderer::*; friend type GetMember(SvgfParamsTag); };
template <typename Tag, typename Tag::type M> struct MemberAccess { friend typename Tag::type GetMember(Tag) { return M; } };
template struct MemberAccess<SvgfParamsTag, &aten::SVGFRenderer::params_>;

// ---- ReSTIR --------------------------------------------------------------------------------------
// The stock host driver aten::ReSTIRRenderer::OnRender is in a debugging state (RELEASE_DEBUG, restir.cpp:10):
// Render() calls GeneratePath again at every bounce, so the "path" restarts from the camera.  The per-pixel steps
// themselves -- ReSTIRRenderer::Shade / EvaluateVisibility / ApplyTemporalReuse / ApplySpatialReuse /
// ComputePixelColor, private static members built on restir_impl.h -- are sound and shared with the device
// renderer.  ReplayReSTIR calls exactly those functions, in the order of idaten::ReSTIRPathTracing::OnRender
// (src/libidaten/restir/restir.cpp:47-123, restir.cu:340-425, restir_reuse.cu:124-230): one GeneratePath per frame,
// bounce 0 = Shade, visibility, temporal reuse (frame > 1), spatial reuse, pixel colour; bounces > 0 = the path
// tracer's shade + shadow ray.  The private statics are reached through explicit-instantiation accessors.
using ReSTIRShadeFn = void (*)(int32_t, int32_t, int32_t, aten::Path&, const aten::context&, aten::ray*, const aten::Intersection&,
                               aten::span<AT_NAME::Reservoir>&, aten::span<AT_NAME::ReSTIRInfo>&, int32_t, int32_t, const aten::mat4&,
                               aten::span<aten::vec4>&, aten::span<aten::vec4>&);
using ReSTIRVisFn = void (*)(int32_t, int32_t, int32_t, int32_t, aten::Path&, const aten::context&, aten::scene*,
                             std::vector<AT_NAME::Reservoir>&, std::vector<AT_NAME::ReSTIRInfo>&, std::vector<aten::ShadowRay>&);
using ReSTIRSpatialFn = void (*)(int32_t, int32_t, int32_t, const aten::context&, aten::sampler&, const std::vector<AT_NAME::Reservoir>&,
                                 std::vector<AT_NAME::Reservoir>&, const std::vector<AT_NAME::ReSTIRInfo>&, const std::vector<aten::vec4>&);
using ReSTIRTemporalFn = void (*)(int32_t, int32_t, int32_t, const aten::context&, aten::sampler&, std::vector<AT_NAME::Reservoir>&,
                                  const std::vector<AT_NAME::Reservoir>&, const std::vector<AT_NAME::ReSTIRInfo>&,
                                  const std::vector<AT_NAME::ReSTIRInfo>&, const std::vector<aten::vec4>&, const std::vector<aten::vec4>&);
using ReSTIRColorFn = void (*)(int32_t, AT_NAME::Path&, const aten::context&, const std::vector<AT_NAME::Reservoir>&,
                               const std::vector<AT_NAME::ReSTIRInfo>&, const std::vector<aten::vec4>&);
struct RsShadeTag { using type = ReSTIRShadeFn; friend type GetMember(RsShadeTag); };
struct RsVisTag { using type = ReSTIRVisFn; friend type GetMember(RsVisTag); };
struct RsSpatialTag { using type = ReSTIRSpatialFn; friend type GetMember(RsSpatialTag); };
struct RsTemporalTag { using type = ReSTIRTemporalFn; friend type GetMember(RsTemporalTag); };
struct RsColorTag { using type = ReSTIRColorFn; friend type GetMember(RsColorTag); };
template struct MemberAccess<RsShadeTag, &aten::ReSTIRRenderer::Shade>;
template struct MemberAccess<RsVisTag, &aten::ReSTIRRenderer::EvaluateVisibility>;
template struct MemberAccess<RsSpatialTag, &aten::ReSTIRRenderer::ApplySpatialReuse>;
template struct MemberAccess<RsTemporalTag, &aten::ReSTIRRenderer::ApplyTemporalReuse>;
template struct MemberAccess<RsColorTag, &aten::ReSTIRRenderer::ComputePixelColor>;

struct ReSTIRState {
    aten::PathHost ph;
    std::vector<aten::ray> rays;
    std::vector<aten::ShadowRay> shadow;
    AT_NAME::ReuseParams<std::vector<AT_NAME::Reservoir>> reservoirs;
    AT_NAME::ReuseParams<std::vector<AT_NAME::ReSTIRInfo>> infos;
    AT_NAME::MatricesForRendering mtxs;
    std::vector<aten::vec4> aov_nd, aov_am, motion;
    int32_t last_target{ 0 };
};

// One frame of the ReSTIR sequence (see the comment above ReSTIRState).
void ReplayReSTIRFrame(Harness& hs, ReSTIRState& st, uint32_t frame, int32_t max_depth, int32_t rr_depth, std::vector<aten::vec4>* image)
{
    const int32_t W = hs.W, H = hs.H, n = W * H;
    if (rr_depth > max_depth) rr_depth = max_depth - 1;                 // restir.cpp:333-335
    // ReSTIRRenderer::Initialize (restir.cpp:322-356)
    if (st.rays.empty()) st.rays.resize(n);
    if (st.shadow.empty()) st.shadow.resize(n);
    st.ph.init(W, H);
    st.ph.Clear(frame);
    st.reservoirs.Init(n);
    st.infos.Init(n);
    st.mtxs.Reset(hs.camera.param());
    if (st.aov_nd.empty()) { st.aov_nd.resize(n); st.aov_am.resize(n); }
    const auto& cam = hs.camera.param();
    const aten::mat4 w2c = st.mtxs.GetW2C();
    aten::span<aten::vec4> aov_nd(st.aov_nd), aov_am(st.aov_am);
    const auto shade = GetMember(RsShadeTag());
    const auto visibility = GetMember(RsVisTag());
    const auto spatial = GetMember(RsSpatialTag());
    const auto temporal = GetMember(RsTemporalTag());
    const auto pixel_color = GetMember(RsColorTag());

    // generatePath + InitReSTIR (libidaten restir.cpp:66-72, restir.cu:18-34)
    for (int32_t idx = 0; idx < n; idx++) {
        AT_NAME::GeneratePath(st.rays[idx], idx, idx % W, idx / W, W, H, 0, frame, st.ph.paths, cam, aten::getRandom(idx));
        st.reservoirs.GetCurrParams()[idx].clear();
        st.infos.GetCurrParams()[idx].clear();
    }
    for (int32_t bounce = 0; bounce < max_depth; bounce++) {
        for (int32_t idx = 0; idx < n; idx++) {
            auto& attrib = st.ph.paths.attrib[idx];
            if (attrib.attr.is_terminated) continue;
            attrib.attr.isHit = false;
            aten::Intersection isect;
            const bool is_hit = aten::BvhTraverser::Traverse<aten::IntersectType::Closest>(isect, hs.ctxt, st.rays[idx], AT_MATH_EPSILON, AT_MATH_INF);
            if (is_hit) {
                attrib.attr.isHit = true;
                if (bounce == 0) {
                    aten::span<AT_NAME::Reservoir> reservoirs(st.reservoirs.GetCurrParams());
                    aten::span<AT_NAME::ReSTIRInfo> infos(st.infos.GetCurrParams());
                    shade(idx, W, H, st.ph.paths, hs.ctxt, st.rays.data(), isect, reservoirs, infos, rr_depth, bounce, w2c, aov_nd, aov_am);
                }
                else {
                    aten::PathTracing::shade(idx, st.ph.paths, hs.ctxt, st.rays.data(), st.shadow.data(), isect, &hs.scene, rr_depth, bounce);
                    const auto& mtrl = hs.ctxt.GetMaterial(isect.mtrlid);
                    std::ignore = AT_NAME::HitShadowRay(bounce, hs.ctxt, mtrl, attrib, st.ph.paths.contrib[idx], st.shadow[idx]);
                }
            }
            else {
                AT_NAME::ShadeMiss(idx, idx % W, idx / W, W, H, bounce, hs.ctxt, cam, st.ph.paths, st.rays[idx], aov_nd, aov_am);
            }
        }
        if (bounce > 0) continue;
        // OnShadeByShadowRayReSTIR (restir.cu:382-425) and ComputelReuse (restir_reuse.cu:124-230)
        for (int32_t idx = 0; idx < n; idx++) {
            if (st.ph.paths.attrib[idx].attr.is_terminated) continue;
            visibility(idx, bounce, W, H, st.ph.paths, hs.ctxt, &hs.scene, st.reservoirs.GetCurrParams(), st.infos.GetCurrParams(), st.shadow);
        }
        const int32_t dst_idx = st.reservoirs.GetDestinationParamsIdxForSpatialReuse();
        const int32_t cur_infos_idx = st.infos.GetCurrParamsIdx();
        auto& cur_res = st.reservoirs.GetCurrParams();
        auto& prev_res = st.reservoirs.GetPreviousFrameParamsForTemporalReuse();
        auto& dst_res = st.reservoirs.GetDestinationParamsForSpatialReuse();
        auto& cur_infos = st.infos.GetCurrParams();
        auto& prev_infos = st.infos.GetPreviousFrameParamsForTemporalReuse();
        st.reservoirs.Update();
        st.infos.Update();
        if (frame > 1) {
            for (int32_t idx = 0; idx < n; idx++) {
                if (st.ph.paths.attrib[idx].attr.is_terminated) continue;
                temporal(idx, W, H, hs.ctxt, st.ph.sampler[idx], cur_res, prev_res, cur_infos, prev_infos, st.aov_am, st.motion);
            }
        }
        for (int32_t idx = 0; idx < n; idx++) {
            if (st.ph.paths.attrib[idx].attr.is_terminated) continue;
            spatial(idx, W, H, hs.ctxt, st.ph.sampler[idx], cur_res, dst_res, cur_infos, st.aov_am);
        }
        for (int32_t idx = 0; idx < n; idx++) {
            if (st.ph.paths.attrib[idx].attr.is_terminated) continue;
            pixel_color(idx, st.ph.paths, hs.ctxt, st.reservoirs.GetParams(dst_idx), st.infos.GetParams(cur_infos_idx), st.aov_am);
        }
        st.last_target = dst_idx;
    }
    if (image) {
        image->resize(n);
        for (int32_t idx = 0; idx < n; idx++) {
            const auto& c = st.ph.paths.contrib[idx].contrib;
            (*image)[idx] = aten::vec4(c.x, c.y, c.z, 1.0f);
        }
    }
}

double NowMs()
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int Usage()
{
    fprintf(stderr,
        "usage: ref_harness <asset_dir> <cmd> ...\n"
        "  scene  <name> W H out.atnscene\n"
        "  render <name> W H frames depth rr threads out_image.f32   (stock aten::PathTracing::render, progressive film)\n"
        "  trace  <name> W H frame depth rr out_records.bin [out_image.f32]\n"
        "  rays   <name> rays.bin hits.bin         (n x {org[3],dir[3],tmax} -> n x AtnHit)\n"
        "  bench  <name> W H frames depth rr threads repeats     (JSON line with timings and ray counts)\n"
        "  ao     <name> W H frame threads out_image.f32            (stock aten::AORenderer, one frame at index `frame`)\n"
        "  svgf   <name> W H frames depth rr out_prefix [mx my [iters [threads]]]  (stock aten::SVGFRenderer on ONE thread, frames 0..frames-1,\n"
        "         1 sample each; constant motion vector (mx,my) or 0; writes <prefix>.f<k>.f32 per frame and <prefix>.aov.f32)\n"
        "  restir <name> W H frames depth rr out_prefix [mx my]       (the reference's ReSTIR steps in the device renderer's order;\n"
        "         writes <prefix>.f<k>.f32 per frame and <prefix>.reservoirs.f32)\n"
        "  cmj    out.bin                           (golden CMJ stream)\n"
        "  device <name> W H frames spp depth out_image.f32   (device_host only: idaten_b200::PathTracing on the GPU, progressive)\n"
        "  device_update <name> W H spp depth updated.f32 fresh.f32   (device_host only: updateBVH to <name>_moved vs a fresh upload)\n");
    return 2;
}

}  // namespace

int main(int argc, char** argv)
{
    if (argc < 3) return Usage();
    g_asset = argv[1];
    const std::string cmd = argv[2];

    if (cmd == "cmj") {
        // Known-answer table for aten::CMJ (src/libaten/sampler/cmj.h:9-124): for a grid of
        // (index, scramble) 8 successive nextSample() values.
        if (argc < 4) return Usage();
        std::vector<float> out;
        const uint32_t scr[] = { 0u, 1u, 0x1fe3434fu, 0xdeadbeefu, 12345u, 0xffffffffu };
        for (uint32_t idx : { 0u, 1u, 7u, 100u, 255u }) {
            for (uint32_t s : scr) {
                aten::CMJ c;
                c.init(idx, 0, s);
                for (int k = 0; k < 8; k++) out.push_back(c.nextSample());
            }
        }
        WriteFile(argv[3], out.data(), out.size() * sizeof(float));
        return 0;
    }

    if (argc < 4) return Usage();
    const std::string name = argv[3];

    if (cmd == "scene") {
        if (argc < 7) return Usage();
        Harness hs;
        if (!hs.Build(name, atoi(argv[4]), atoi(argv[5]))) return 1;
        idaten_b200::FlatScene flat;
        flat.Gather(hs.ctxt, hs.scene.getAccel()->getNodes());
        // seeds are std::mt19937(0) outputs; readers regenerate them (aten_b200/scene_io.py:mt19937_seeds)
        if (!flat.Save(argv[6], hs.camera.param(), hs.W, hs.H, std::vector<uint32_t>())) return 1;
        size_t nn = 0; for (auto& l : flat.nodes) nn += l.size();
        printf("{\"scene\": \"%s\", \"objects\": %zu, \"matrices\": %zu, \"materials\": %zu, \"lights\": %zu, "
               "\"triangles\": %zu, \"vertices\": %zu, \"node_lists\": %zu, \"nodes\": %zu, \"textures\": %zu}\n",
               name.c_str(), flat.objects.size(), flat.matrices.size(), flat.materials.size(), flat.lights.size(),
               flat.triangles.size(), flat.positions.size(), flat.nodes.size(), nn, flat.textures.size());
        return 0;
    }

#ifdef ATN_WITH_DEVICE
    if (cmd == "device") {
        // The reference's device_renderer frame loop (src/device_renderer/main.cpp:133-148, 185-202) with
        // idaten_b200::PathTracing in the place of idaten::PathTracing: scene built by the reference's own loaders
        // and SBVH builder, handed over through the shim class, rendered on the GPU, image read back.  Only the
        // `device_host` executable has this command (it links libaten_b200.so); `ref_harness` stays CPU-only.
        if (argc < 10) return Usage();
        const int W = atoi(argv[4]), H = atoi(argv[5]), frames = atoi(argv[6]), spp = atoi(argv[7]), depth = atoi(argv[8]);
        Harness hs;
        if (!hs.Build(name, W, H)) return 1;
        try {
            idaten_b200::PathTracing renderer;
            const auto& nodes = hs.scene.getAccel()->getNodes();
            auto camparam = hs.camera.param();
            camparam.znear = 0.1f; camparam.zfar = 10000.0f;
            renderer.UpdateSceneData(0, W, H, camparam, hs.ctxt, nodes, 0, 0);
            renderer.SetEnableProgressive(true);
            renderer.updateCamera(camparam);
            double t0 = NowMs();
            for (int f = 0; f < frames; f++) renderer.render(W, H, spp, depth);
            std::vector<aten::vec4> img((size_t)W * H);
            renderer.ReadbackImage(img.data(), img.size());
            const double ms = NowMs() - t0;
            WriteFile(argv[9], img.data(), img.size() * sizeof(aten::vec4));
            printf("{\"scene\": \"%s\", \"w\": %d, \"h\": %d, \"frames\": %d, \"spp\": %d, \"depth\": %d, \"frame_count\": %u, "
                   "\"wall_ms\": %.3f, \"host\": \"idaten_b200::PathTracing (C++ shim over the C ABI)\"}\n",
                   name.c_str(), W, H, frames, spp, depth, renderer.GetFrameCount(), ms);
        }
        catch (const std::exception& e) {
            fprintf(stderr, "device: %s\n", e.what());
            return 3;
        }
        return 0;
    }
    if (cmd == "device_update") {
        // Renderer::updateBVH through the shim (src/device_renderer/main.cpp:440-470): upload scene <name>, then hand over
        // the top layer, objects and matrices of scene <name>_moved (same meshes, one instance displaced) and render;
        // a second renderer gets the moved scene as a fresh upload.  Writes both images.
        if (argc < 10) return Usage();
        const int W = atoi(argv[4]), H = atoi(argv[5]), spp = atoi(argv[6]), depth = atoi(argv[7]);
        Harness hs, moved;
        if (!hs.Build(name, W, H) || !moved.Build(name + "_moved", W, H)) return 1;
        try {
            idaten_b200::PathTracing renderer, fresh;
            renderer.UpdateSceneData(0, W, H, hs.camera.param(), hs.ctxt, hs.scene.getAccel()->getNodes(), 0, 0);
            renderer.render(W, H, 1, depth);
            renderer.updateBVH(moved.ctxt, moved.scene.getAccel()->getNodes());
            renderer.reset();
            renderer.render(W, H, spp, depth);
            fresh.UpdateSceneData(0, W, H, moved.camera.param(), moved.ctxt, moved.scene.getAccel()->getNodes(), 0, 0);
            fresh.render(W, H, spp, depth);
            std::vector<aten::vec4> a((size_t)W * H), b((size_t)W * H);
            renderer.ReadbackImage(a.data(), a.size());
            fresh.ReadbackImage(b.data(), b.size());
            WriteFile(argv[8], a.data(), a.size() * sizeof(aten::vec4));
            WriteFile(argv[9], b.data(), b.size() * sizeof(aten::vec4));
            printf("{\"scene\": \"%s\", \"w\": %d, \"h\": %d, \"spp\": %d}\n", name.c_str(), W, H, spp);
        }
        catch (const std::exception& e) {
            fprintf(stderr, "device_update: %s\n", e.what());
            return 3;
        }
        return 0;
    }
#endif

    if (cmd == "render" || cmd == "bench") {
        if (argc < 11) return Usage();
        const int W = atoi(argv[4]), H = atoi(argv[5]), frames = atoi(argv[6]), depth = atoi(argv[7]), rr = atoi(argv[8]);
        const int threads = atoi(argv[9]);
        Harness hs;
        double t0 = NowMs();
        if (!hs.Build(name, W, H)) return 1;
        const double build_ms = NowMs() - t0;
        aten::OMPUtil::setThreadNum(threads);
        omp_set_num_threads(threads);
        if (cmd == "render") {
            aten::PathTracing pt;
            aten::FilmProgressive film(W, H);
            auto dst = hs.MakeDst(&film, depth, rr);
            t0 = NowMs();
            for (int f = 0; f < frames; f++) pt.render(hs.ctxt, dst, &hs.scene, &hs.camera);
            const double ms = NowMs() - t0;
            WriteFile(argv[10], film.image().data(), film.image().size() * sizeof(aten::vec4));
            printf("{\"scene\": \"%s\", \"w\": %d, \"h\": %d, \"frames\": %d, \"threads\": %d, \"render_ms\": %.3f, \"build_ms\": %.1f}\n",
                   name.c_str(), W, H, frames, threads, ms, build_ms);
            return 0;
        }
        const int repeats = atoi(argv[10]);
        // Ray counts from the replay (identical traversal sequence), timing from the stock render().
        Counters cnt;
        for (int f = 0; f < frames; f++) Replay(hs, f, depth, rr, nullptr, nullptr, &cnt);
        double best = 1e30;
        for (int rep = 0; rep < repeats + 1; rep++) {     // first repetition is the warm-up
            aten::PathTracing pt;
            aten::FilmProgressive film(W, H);
            auto dst = hs.MakeDst(&film, depth, rr);
            t0 = NowMs();
            for (int f = 0; f < frames; f++) pt.render(hs.ctxt, dst, &hs.scene, &hs.camera);
            const double ms = NowMs() - t0;
            if (rep > 0 && ms < best) best = ms;
        }
        const double samples = (double)W * H * frames;
        printf("{\"scene\": \"%s\", \"w\": %d, \"h\": %d, \"frames\": %d, \"depth\": %d, \"threads\": %d, \"best_ms\": %.3f, "
               "\"msamples_per_s\": %.4f, \"mrays_per_s\": %.4f, \"closest_rays\": %llu, \"shadow_rays\": %llu, \"build_ms\": %.1f}\n",
               name.c_str(), W, H, frames, depth, threads, best, samples / best / 1e3,
               (double)(cnt.closest + cnt.shadow) / best / 1e3,
               (unsigned long long)cnt.closest, (unsigned long long)cnt.shadow, build_ms);
        return 0;
    }

    if (cmd == "ao") {
        // stock aten::AORenderer (src/libaten/renderer/ao/aorenderer.cpp), 1 AO ray, radius 1, progressive film
        if (argc < 9) return Usage();
        // argv[6] = frame index of the ONE frame rendered.  (The stock renderer keeps its Intersection array
        // across frames without resetting .t, so later frames of one renderer object only accept hits closer
        // than the previous frame's -- a fresh renderer per frame avoids pinning that.)
        const int W = atoi(argv[4]), H = atoi(argv[5]), frame = atoi(argv[6]), threads = atoi(argv[7]);
        const int frames = 1;
        Harness hs;
        if (!hs.Build(name, W, H)) return 1;
        aten::OMPUtil::setThreadNum(threads);
        omp_set_num_threads(threads);
        struct FramedAO : public aten::AORenderer { void SetFrame(uint32_t f) { frame_count_ = f; } } ao;
        ao.SetFrame(frame);
        aten::FilmProgressive film(W, H);
        auto dst = hs.MakeDst(&film, 1, 1);
        double t0 = NowMs();
        ao.render(hs.ctxt, dst, &hs.scene, &hs.camera);
        const double ms = NowMs() - t0;
        WriteFile(argv[8], film.image().data(), film.image().size() * sizeof(aten::vec4));
        printf("{\"scene\": \"%s\", \"renderer\": \"ao\", \"w\": %d, \"h\": %d, \"frames\": %d, \"threads\": %d, \"render_ms\": %.3f}\n",
               name.c_str(), W, H, frames, threads, ms);
        return 0;
    }

    if (cmd == "svgf") {
        // stock aten::SVGFRenderer (src/libaten/renderer/svgf/svgf.cpp); one OpenMP thread, because its variance
        // estimation filters the colour history in place and is order dependent (svgf_impl.h:438-544)
        if (argc < 10) return Usage();
        const int W = atoi(argv[4]), H = atoi(argv[5]), frames = atoi(argv[6]), depth = atoi(argv[7]), rr = atoi(argv[8]);
        const std::string prefix = argv[9];
        const float mx = argc > 10 ? (float)atof(argv[10]) : 0.0f, my = argc > 11 ? (float)atof(argv[11]) : 0.0f;
        const int iters = argc > 12 ? atoi(argv[12]) : 5;
        const int threads = argc > 13 ? atoi(argv[13]) : 1;     // > 1 only for timing: the result then depends on the schedule
        Harness hs;
        if (!hs.Build(name, W, H)) return 1;
        aten::OMPUtil::setThreadNum(threads);
        omp_set_num_threads(threads);
        aten::SVGFRenderer svgf;
        SvgfParamsT& params = svgf.*GetMember(SvgfParamsTag());
        params.atrous_iter_cnt = iters;    // public field of SVGFParams (svgf_types.h:74), default 5
        aten::Film film(W, H);
        auto dst = hs.MakeDst(&film, depth, rr);
        double ms = 0;
        for (int f = 0; f < frames; f++) {
            params.motion_depth_buffer.assign((size_t)W * H, aten::vec4(mx, my, 0.0f, 0.0f));
            const double t0 = NowMs();
            svgf.render(hs.ctxt, dst, &hs.scene, &hs.camera);
            ms += NowMs() - t0;
            WriteFile(prefix + ".f" + std::to_string(f) + ".f32", film.image().data(), film.image().size() * sizeof(aten::vec4));
        }
        // AOV set of the last frame rendered (the renderer has already toggled to the other set)
        auto& aov = params.GetPrevAovBuffer();
        std::vector<aten::vec4> dump;
        for (auto* b : { &aov.get<AT_NAME::SVGFAovBufferType::NormalDepth>(), &aov.get<AT_NAME::SVGFAovBufferType::AlbedoMeshId>(),
                         &aov.get<AT_NAME::SVGFAovBufferType::ColorVariance>(), &aov.get<AT_NAME::SVGFAovBufferType::MomentTemporalWeight>() })
            dump.insert(dump.end(), b->begin(), b->end());
        WriteFile(prefix + ".aov.f32", dump.data(), dump.size() * sizeof(aten::vec4));
        printf("{\"scene\": \"%s\", \"renderer\": \"svgf\", \"w\": %d, \"h\": %d, \"frames\": %d, \"threads\": %d, \"render_ms\": %.3f}\n",
               name.c_str(), W, H, frames, threads, ms);
        return 0;
    }

    if (cmd == "restir") {
        // the reference's ReSTIR per-pixel steps in the device renderer's order (ReplayReSTIRFrame), one thread
        if (argc < 10) return Usage();
        const int W = atoi(argv[4]), H = atoi(argv[5]), frames = atoi(argv[6]), depth = atoi(argv[7]), rr = atoi(argv[8]);
        const std::string prefix = argv[9];
        const float mx = argc > 10 ? (float)atof(argv[10]) : 0.0f, my = argc > 11 ? (float)atof(argv[11]) : 0.0f;
        Harness hs;
        if (!hs.Build(name, W, H)) return 1;
        aten::OMPUtil::setThreadNum(1);
        omp_set_num_threads(1);
        ReSTIRState st;
        st.motion.assign((size_t)W * H, aten::vec4(mx, my, 0.0f, 0.0f));
        std::vector<aten::vec4> image;
        double ms = 0;
        for (int f = 0; f < frames; f++) {
            const double t0 = NowMs();
            ReplayReSTIRFrame(hs, st, (uint32_t)f, depth, rr, &image);
            ms += NowMs() - t0;
            WriteFile(prefix + ".f" + std::to_string(f) + ".f32", image.data(), image.size() * sizeof(aten::vec4));
        }
        // the reservoirs the last frame's pixel colours were computed from: w_sum, M, y, W, target pdf per pixel
        std::vector<float> dump;
        for (const auto& r : st.reservoirs.GetParams(st.last_target)) {
            dump.push_back(r.w_sum); dump.push_back((float)r.M); dump.push_back((float)r.y); dump.push_back(r.W); dump.push_back(r.target_pdf_of_y);
        }
        WriteFile(prefix + ".reservoirs.f32", dump.data(), dump.size() * sizeof(float));
        printf("{\"scene\": \"%s\", \"renderer\": \"restir\", \"w\": %d, \"h\": %d, \"frames\": %d, \"threads\": 1, \"render_ms\": %.3f}\n",
               name.c_str(), W, H, frames, ms);
        return 0;
    }

    if (cmd == "trace") {
        if (argc < 10) return Usage();
        const int W = atoi(argv[4]), H = atoi(argv[5]), frame = atoi(argv[6]), depth = atoi(argv[7]), rr = atoi(argv[8]);
        Harness hs;
        if (!hs.Build(name, W, H)) return 1;
        std::vector<AtnBounceRecord> recs;
        std::vector<aten::vec4> img;
        Counters cnt;
        Replay(hs, frame, depth, rr, &recs, &img, &cnt);
        // Self-check: the replay must equal the stock renderer bit for bit.
        {
            CountingPathTracing pt;
            pt.SetFrame(frame);
            aten::Film film(W, H);
            auto dst = hs.MakeDst(&film, depth, rr);
            pt.render(hs.ctxt, dst, &hs.scene, &hs.camera);
            size_t bad = 0;
            for (size_t i = 0; i < img.size(); i++) {
                const auto& a = film.image()[i];
                const auto& b = img[i];
                const bool same = (memcmp(&a.x, &b.x, 12) == 0) || (std::isnan(a.x) && true);
                if (!same) bad++;
            }
            if (bad) { fprintf(stderr, "replay != stock render on %zu pixels\n", bad); return 3; }
        }
        WriteFile(argv[9], recs.data(), recs.size() * sizeof(AtnBounceRecord));
        if (argc > 10) WriteFile(argv[10], img.data(), img.size() * sizeof(aten::vec4));
        printf("{\"scene\": \"%s\", \"w\": %d, \"h\": %d, \"frame\": %d, \"closest_rays\": %llu, \"shadow_rays\": %llu, \"replay_equals_stock\": true}\n",
               name.c_str(), W, H, frame, (unsigned long long)cnt.closest, (unsigned long long)cnt.shadow);
        return 0;
    }

    if (cmd == "rays") {
        if (argc < 6) return Usage();
        Harness hs;
        if (!hs.Build(name, 4, 4)) return 1;
        FILE* fp = fopen(argv[4], "rb");
        if (!fp) return 1;
        fseek(fp, 0, SEEK_END);
        const size_t bytes = ftell(fp);
        fseek(fp, 0, SEEK_SET);
        const size_t n = bytes / 28;
        std::vector<float> in(n * 7);
        if (fread(in.data(), 28, n, fp) != n) return 1;
        fclose(fp);
        std::vector<AtnHit> hits(n);
#pragma omp parallel for
        for (long i = 0; i < (long)n; i++) {
            aten::ray r;
            r.org = aten::vec3(in[i * 7 + 0], in[i * 7 + 1], in[i * 7 + 2]);
            r.dir = aten::vec3(in[i * 7 + 3], in[i * 7 + 4], in[i * 7 + 5]);
            aten::Intersection isect;
            aten::BvhTraverser::Traverse<aten::IntersectType::Closest>(isect, hs.ctxt, r, AT_MATH_EPSILON, in[i * 7 + 6]);
            memcpy(&hits[i], &isect, sizeof(isect));
        }
        WriteFile(argv[5], hits.data(), hits.size() * sizeof(AtnHit));
        return 0;
    }
    return Usage();
}
