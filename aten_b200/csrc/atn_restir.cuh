// atn_restir.cuh -- the per-pixel steps of the ReSTIR renderer (N1), one text for the sm_100a kernels
// (atn_kernels.cu) and for the CPU restatement under oracle/, which compiles this header with g++.
//
// Reference: src/libaten/renderer/restir/restir_impl.h (shared by aten::ReSTIRRenderer and idaten::ReSTIRPathTracing),
// restir_types.h, and ReSTIRRenderer::Shade (src/libaten/renderer/restir/restir.cpp:100-241).  Order of the steps:
// idaten::ReSTIRPathTracing::OnRender (src/libidaten/restir/restir.cpp:47-123, restir.cu:340-425,
// restir_reuse.cu:124-230).
#pragma once

#include "atn_math.cuh"

namespace atn {

// AT_NAME::Reservoir (restir_types.h:6-57)
struct Reservoir {
    float w_sum;          // sum of weights
    int32_t M;            // samples seen so far
    int32_t y;            // the output sample (light index), -1: none
    float W;              // stochastic weight of y
    float target_pdf;     // target pdf of y
    LightSample ls;       // light_sample_

    ATN_HD void clear() { w_sum = 0.0f; M = 0; y = -1; target_pdf = 0.0f; W = 0.0f; }
    ATN_HD bool valid() const { return y >= 0; }
    ATN_HD bool update(const LightSample& s, int32_t sample, float weight, int32_t m, float u)
    {
        w_sum += weight;
        const bool accepted = u < weight / w_sum;
        if (accepted) { ls = s; y = sample; }
        M += m;
        return accepted;
    }
};

// AT_NAME::ReSTIRInfo (restir_types.h:61-99)
struct RestirInfo {
    V3 nml; int32_t mtrl_idx;
    V3 wi; float u;
    V3 point_to_light; float v;
    V3 p; float pre_sampled_r;
    V3 throughput; int32_t mesh_id;

    ATN_HD void clear()
    {
        nml = V3(0.0f); mtrl_idx = -1; wi = V3(0.0f); u = 0.0f; v = 0.0f; pre_sampled_r = 0.0f; mesh_id = -1;
    }
};

// restir::_detail::ComputeRadiance (restir_impl.h:18-52)
ATN_HD V3 restir_radiance(const SceneView& sc, const LightSample& ls, uint32_t light_attrib, const V3& normal, const V3& ray_dir,
                          const Mtrl& mtrl, float u, float v)
{
    const V3 nml_light = ls.nml;
    const V3 dir_to_light = ls.dir;
    const float cos_shadow = fabsf(dot(normal, dir_to_light));
    const float cos_light = fabsf(dot(nml_light, -dir_to_light));
    const float dist2 = ls.dist * ls.dist;
    const MtrlSample ev = material_bsdf(sc, mtrl, normal, ray_dir, dir_to_light, u, v);
    const bool singular = (light_attrib & 1u) != 0, infinite = (light_attrib & 2u) != 0;
    const float geometry_term = (singular || infinite) ? cos_shadow * cos_light : cos_shadow * cos_light / dist2;
    return ev.bsdf * ls.color * geometry_term;
}

// restir::_detail::ComputeTargetPDF (restir_impl.h:54-82)
ATN_HD float restir_target_pdf(const SceneView& sc, const LightSample& ls, uint32_t light_attrib, const V3& normal, const V3& ray_dir,
                               const Mtrl& mtrl, float u, float v)
{
    const float pdf = material_pdf(sc, mtrl, normal, ray_dir, ls.dir, u, v);
    if (pdf == 0.0f) return 0.0f;
    const V3 energy = restir_radiance(sc, ls, light_attrib, normal, ray_dir, mtrl, u, v);
    return (energy.x + energy.y + energy.z) / 3;
}

ATN_HD bool restir_isfinite(float f) { return !(f != f) && f <= kInf && f >= -kInf; }

// restir::GenerateInitialCandidate<CONTEXT, 1> (restir_impl.h:84-161): the host renderer and this path draw ONE
// light candidate per pixel (MaxLightCount = 1, restir.cpp:203).
ATN_HD int32_t restir_initial_candidate(Reservoir* res, const Mtrl& mtrl, const SceneView& sc, const V3& org, const V3& normal,
                                        const V3& ray_dir, float u, float v, Sampler* sampler)
{
    const int32_t max_light_num = sc.light_count;
    const int32_t light_cnt = 1 < max_light_num ? 1 : max_light_num;
    res->clear();
    float candidate_target_pdf = 0.0f;
    const float light_select_prob = 1.0f / max_light_num;
    for (int32_t i = 0; i < light_cnt; i++) {
        const float r_light = sampler->next();
        int32_t light_pos = (int32_t)(r_light * max_light_num);
        light_pos = light_pos < 0 ? 0 : (light_pos > max_light_num - 1 ? max_light_num - 1 : light_pos);
        const AtnLightParameter& light = sc.lights[light_pos];
        LightSample ls;
        light_sample(&ls, light, sc, org, normal, sampler);
        const float sampling_pdf = ls.pdf * light_select_prob;
        const float target_pdf = restir_target_pdf(sc, ls, light.attrib, normal, ray_dir, mtrl, u, v);
        const float weight = sampling_pdf > 0 ? target_pdf / sampling_pdf : 0.0f;
        const float r = sampler->next();
        if (res->update(ls, light_pos, weight, 1, r)) candidate_target_pdf = target_pdf;
    }
    if (candidate_target_pdf > 0.0f) {
        res->target_pdf = candidate_target_pdf;
        res->W = res->w_sum / (res->target_pdf * res->M);
    }
    if (!restir_isfinite(res->W)) res->clear();
    return res->y;
}

// First half of restir::EvaluateVisibility (restir_impl.h:163-199): the shadow ray towards the reservoir's sample.
// The caller traces it (HitShadowRay: closest hit with t_max = dist - 1e-9, scene::hitLight) and passes the answer
// to restir_visibility_result.  Note what the reference leaves untouched: ShadowRay::lightcontrib keeps whatever the
// last path-traced NEE of this pixel stored, and HitShadowRay adds it to the contribution when the light is reached.
ATN_HD void restir_shadow_ray(const Reservoir& res, RestirInfo* info, V3* org, V3* dir, float* dist)
{
    *org = info->p + kEps * info->nml;
    V3 d = res.ls.pos - *org;
    info->point_to_light = res.ls.pos - *org;
    *dist = length(d);
    *dir = d / *dist;
}
ATN_HD void restir_visibility_result(Reservoir* res, bool reached)
{
    if (!reached) { res->w_sum = 0.0f; res->W = 0.0f; res->target_pdf = 0.0f; res->y = -1; }     // M is kept (restir_impl.h:201-210)
}

// restir::_detail::IsAcceptableNeighbor (restir_impl.h:212-228).  Only the TYPE of the neighbour's material takes part,
// so the neighbour's 248-byte MaterialParameter is not copied (FillMaterial of an invalid id yields a Diffuse material).
ATN_HD int32_t restir_material_type(const SceneView& sc, int32_t mtrl_idx)
{
    return mtrl_idx >= 0 ? sc.materials[mtrl_idx].type : (int32_t)ATN_MTRL_DIFFUSE;
}
ATN_HD bool restir_acceptable(const Mtrl& m, int32_t mesh_id, const V3& n, int32_t n_type, int32_t n_mesh_id, const V3& nn)
{
    return ((int32_t)m.type == n_type) && (mesh_id == n_mesh_id) && (dot(n, nn) >= 0.95f);
}

// restir::ApplyTemporalReuse (restir_impl.h:230-346).  `motion` = the centre pixel's motion vector (xy).
ATN_HD void restir_temporal(int32_t idx, int32_t width, int32_t height, const SceneView& sc, Sampler* sampler, Reservoir* combined,
                            const RestirInfo& self, const Reservoir* prev_res, const RestirInfo* prev_infos,
                            const AtnVec4* aov_albedo_meshid, const AtnVec4* motion_depth)
{
    const int32_t ix = idx % width, iy = idx / width;
    const Mtrl mtrl = fill_material(sc, self.mtrl_idx);
    const V3 normal = self.nml;
    const int32_t mesh_id = (int32_t)aov_albedo_meshid[idx].w;
    float candidate_target_pdf = combined->valid() ? combined->target_pdf : 0.0f;
    const int32_t maxM = 20 * combined->M;
    const AtnVec4 md = motion_depth ? motion_depth[idx] : v4(0.0f);
    const int32_t px = (int32_t)(ix + md.x * width);
    const int32_t py = (int32_t)(iy + md.y * height);
    bool acceptable = (0 <= px && px <= width - 1) && (0 <= py && py <= height - 1);
    if (acceptable) {
        LightSample ls;
        ls.pos = V3(0.0f); ls.dir = V3(0.0f); ls.dist = 0.0f; ls.nml = V3(0.0f); ls.color = V3(0.0f); ls.pdf = 0.0f; ls.attrib = 0;
        const int32_t nidx = py * width + px;
        const Reservoir& nr = prev_res[nidx];
        const int32_t m = nr.M < maxM ? nr.M : maxM;
        if (nr.valid()) {
            const RestirInfo& ni = prev_infos[nidx];
            const bool valid_mtrl = ni.mtrl_idx >= 0;
            acceptable = valid_mtrl && restir_acceptable(mtrl, mesh_id, normal, restir_material_type(sc, ni.mtrl_idx), ni.mesh_id, ni.nml);
            if (acceptable) {
                const int32_t light_pos = nr.y;
                const AtnLightParameter& light = sc.lights[light_pos];
                light_sample(&ls, light, sc, self.p, ni.nml, sampler);
                const float target_pdf = restir_target_pdf(sc, ls, light.attrib, self.nml, self.wi, mtrl, self.u, self.v);
                const float weight = target_pdf * nr.W * m;
                const float r = sampler->next();
                if (combined->update(ls, light_pos, weight, m, r)) candidate_target_pdf = target_pdf;
            }
        }
        else {
            combined->update(ls, -1, 0.0f, m, 0.0f);
        }
    }
    if (candidate_target_pdf > 0.0f) {
        combined->target_pdf = candidate_target_pdf;
        combined->W = combined->w_sum / (combined->target_pdf * combined->M);
    }
    if (!restir_isfinite(combined->W)) combined->clear();
}

// restir::ApplySpatialReuse (restir_impl.h:348-460): the 3x3 neighbourhood in row-major order, into a separate
// destination reservoir.
ATN_HD void restir_spatial(int32_t idx, int32_t width, int32_t height, const SceneView& sc, Sampler* sampler, Reservoir* combined,
                           const Reservoir* reservoirs, const RestirInfo* infos, const AtnVec4* aov_albedo_meshid)
{
    const int32_t ix = idx % width, iy = idx / width;
    const RestirInfo& self = infos[idx];
    const Mtrl mtrl = fill_material(sc, self.mtrl_idx);
    const V3 normal = self.nml;
    const int32_t mesh_id = (int32_t)aov_albedo_meshid[idx].w;
    combined->clear();
    float candidate_target_pdf = 0.0f;
    int32_t M_sum = 0;
    for (int32_t oy = -1; oy <= 1; oy++) {
        for (int32_t ox = -1; ox <= 1; ox++) {
            const int32_t xx = ix + ox, yy = iy + oy;
            if (!((0 <= xx && xx <= width - 1) && (0 <= yy && yy <= height - 1))) continue;
            const int32_t nidx = yy * width + xx;
            const Reservoir& nr = reservoirs[nidx];
            M_sum += nr.M;
            if (!nr.valid()) continue;
            const RestirInfo& ni = infos[nidx];
            const bool valid_mtrl = ni.mtrl_idx >= 0;
            const int32_t n_mesh_id = (int32_t)aov_albedo_meshid[nidx].w;
            if (!(valid_mtrl && restir_acceptable(mtrl, mesh_id, normal, restir_material_type(sc, ni.mtrl_idx), n_mesh_id, ni.nml))) continue;
            const int32_t light_pos = nr.y;
            const AtnLightParameter& light = sc.lights[light_pos];
            LightSample ls;
            light_sample(&ls, light, sc, self.p, ni.nml, sampler);
            const float target_pdf = restir_target_pdf(sc, ls, light.attrib, self.nml, self.wi, mtrl, self.u, self.v);
            const int32_t m = nr.M;
            const float weight = target_pdf * nr.W * m;
            const float r = sampler->next();
            if (combined->update(ls, light_pos, weight, m, r)) candidate_target_pdf = target_pdf;
        }
    }
    combined->M = M_sum;
    if (candidate_target_pdf > 0.0f) {
        combined->target_pdf = candidate_target_pdf;
        combined->W = combined->w_sum / (combined->target_pdf * combined->M);
    }
    if (!restir_isfinite(combined->W)) combined->clear();
}

// restir::ComputePixelColor (restir_impl.h:462-503) times the throughput kept in the info
// (ReSTIRRenderer::ComputePixelColor, restir.cpp:358-384).  Returns false when the reservoir holds no sample.
ATN_HD bool restir_pixel_color(const SceneView& sc, const Reservoir& res, const RestirInfo& info, const AtnVec4& albedo_meshid, V3* result)
{
    if (!res.valid()) return false;
    const Mtrl mtrl = fill_material(sc, info.mtrl_idx);
    const AtnLightParameter& light = sc.lights[res.y];
    const V3 le = restir_radiance(sc, res.ls, light.attrib, info.nml, info.wi, mtrl, info.u, info.v);
    V3 contrib = le * res.W;
    contrib = contrib * V3(albedo_meshid.x, albedo_meshid.y, albedo_meshid.z);
    *result = contrib * info.throughput;
    return true;
}

// ReSTIRRenderer::Shade up to the candidate generation (restir.cpp:100-213): hit evaluation, material, albedo (white by
// default), normal map BEFORE the flip, flip unless translucent or emissive, ReSTIRInfo, AOVs of bounce 0 (shading
// normal, clip-space w, mesh id), implicit light.  Returns false when the path ended on a light.  The Surface record is
// what the BSDF stage (shade_bsdf) continues from, as on the path-tracing path.
ATN_HD bool restir_surface(const SceneView& sc, const Ray& ray, int32_t objid, int32_t primid, float a, float b, float t,
                           int32_t mtrlid, int32_t meshid, int32_t bounce, const AtnMat4& w2c, PathState* st, Surface* sf,
                           RestirInfo* info, AtnVec4* aov_normal_depth, AtnVec4* aov_albedo_meshid)
{
    const AtnObjectParameter obj = sc.objects[objid];
    HitRecord rec;
    evaluate_hit(sc, obj, ray, primid, a, b, t, mtrlid, &rec);
    const bool backfacing = dot(rec.normal, -ray.dir) < 0.0f;
    V3 orienting = rec.normal;
    const Mtrl mtrl = fill_material(sc, rec.mtrlid);
    const AtnVec4 albedo4 = sample_texture(sc, mtrl.albedoMap, rec.u, rec.v, v4(1.0f));
    orienting = apply_normal_map(sc, mtrl.normalMap, orienting, rec.u, rec.v);
    if (!mtrl.is_translucent() && !mtrl.is_emissive() && backfacing) orienting = -orienting;

    info->clear();
    info->nml = orienting; info->mtrl_idx = (int32_t)(int16_t)rec.mtrlid;
    info->wi = ray.dir; info->u = rec.u; info->v = rec.v; info->p = rec.p;
    info->pre_sampled_r = -1.0f;            // material::applyNormal returns -1 for everything but car paint
    info->mesh_id = meshid;
    info->throughput = st->throughput;

    if (bounce == 0 && aov_normal_depth && aov_albedo_meshid)
        fill_basic_aovs(aov_normal_depth, orienting, rec.p, w2c, aov_albedo_meshid, albedo4, meshid);

    // HitImplicitLight (pathtracing_impl.h:395-452)
    if (mtrl.is_emissive() && !backfacing && obj.light_id >= 0) {
        const AtnLightParameter& light = sc.lights[obj.light_id];
        const float lum = light.scale * light.intensity / rec.area;
        const V3 light_color = V3(light.light_color) * lum;
        float weight = 1.0f;
        if (bounce > 0) {
            const float cosLight = dot(rec.normal, -ray.dir);
            const float dist2 = squared_length(rec.p - ray.org);
            if (cosLight >= 0) {
                float pdfLight = 1 / rec.area;
                pdfLight = pdfLight * dist2 / cosLight;
                weight = st->pdfb / (st->pdfb + pdfLight);
            }
        }
        st->contrib = st->contrib + st->throughput * weight * light_color;
        st->terminated = true;
        return false;
    }
    sf->p = rec.p; sf->u = rec.u; sf->n = orienting; sf->v = rec.v;
    sf->albedo = V3(albedo4); sf->mtrlid = rec.mtrlid;
    return true;
}

}  // namespace atn
