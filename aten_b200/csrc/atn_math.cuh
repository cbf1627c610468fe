// atn_math.cuh -- scalar shading / sampling math of the path-tracing hot path, written once as
// host+device inline functions.  The CUDA kernels (atn_kernels.cu) compile it for sm_100a; the CPU
// restatement under oracle/ compiles the same text with g++ so that the arithmetic can be pinned,
// bit for bit, against the reference's own CPU renderer (oracle/_ref) without a GPU.
//
// Every function cites the reference code it restates (paths relative to the reference root).
// The operation ORDER of every floating-point expression follows the reference (glm's
// dot = (x*x + y*y) + z*z, normalize = v * (1/sqrt(dot)), cross = a.y*b.z - b.y*a.z ...), because
// parity is judged bit-exact on the traversal and to float tolerance on shading.  The file must be
// compiled without FMA contraction (nvcc -fmad=false; x86-64 g++ has no FMA by default).
#pragma once

#include <stdint.h>
#include <math.h>
#include <string.h>

#include "../../include/aten_b200.h"

#if defined(__CUDACC__)
#define ATN_HD __host__ __device__ __forceinline__
// ATN_NOINLINE_LEVEL (kernel tuning): 0 = everything inlined; 1 = the rarely taken BSDFs (Disney,
// Beckmann, refraction) are real functions; 2 = also the sampler, sin/cos users, hit evaluation and
// light sampling.  The shade kernel is instruction-fetch bound (ncu: "no_instruction" is its top
// stall), so code size matters more than call overhead there.
#ifndef ATN_NOINLINE_LEVEL
#define ATN_NOINLINE_LEVEL 0
#endif
#if ATN_NOINLINE_LEVEL >= 1
#define ATN_HD_COLD __host__ __device__ __noinline__
#else
#define ATN_HD_COLD ATN_HD
#endif
#if ATN_NOINLINE_LEVEL >= 2
#define ATN_HD_CALL __host__ __device__ __noinline__
#else
#define ATN_HD_CALL ATN_HD
#endif
#else
#define ATN_HD inline
#define ATN_HD_COLD inline
#define ATN_HD_CALL inline
#endif

namespace atn {

constexpr float kPi = 3.14159265358979323846f;       // AT_MATH_PI, src/libaten/math/math.h:10
constexpr float kPi2 = kPi * 2;                       // AT_MATH_PI_2
constexpr float kPiHalf = kPi * 0.5f;
constexpr float kInf = 3.402823466e+38f;              // AT_MATH_INF = FLT_MAX, math.h:14
constexpr float kEps = 1e-9f;                         // AT_MATH_EPSILON, math.h:15

// ---------------------------------------------------------------------------------------------
// vec3 with glm's evaluation order (3rdparty/glm/glm/detail/func_geometric.inl:48-94)
// ---------------------------------------------------------------------------------------------
struct V3 {
    float x, y, z;
    ATN_HD V3() : x(0), y(0), z(0) {}
    ATN_HD V3(float a) : x(a), y(a), z(a) {}
    ATN_HD V3(float a, float b, float c) : x(a), y(b), z(c) {}
    ATN_HD V3(const AtnVec3& v) : x(v.x), y(v.y), z(v.z) {}
    ATN_HD V3(const AtnVec4& v) : x(v.x), y(v.y), z(v.z) {}
};
ATN_HD V3 operator+(const V3& a, const V3& b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
ATN_HD V3 operator-(const V3& a, const V3& b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
ATN_HD V3 operator*(const V3& a, const V3& b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
ATN_HD V3 operator/(const V3& a, const V3& b) { return V3(a.x / b.x, a.y / b.y, a.z / b.z); }
ATN_HD V3 operator*(const V3& a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
ATN_HD V3 operator*(float s, const V3& a) { return V3(s * a.x, s * a.y, s * a.z); }
ATN_HD V3 operator/(const V3& a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
ATN_HD V3 operator+(const V3& a, float s) { return V3(a.x + s, a.y + s, a.z + s); }
ATN_HD V3 operator-(const V3& a, float s) { return V3(a.x - s, a.y - s, a.z - s); }
ATN_HD V3 operator-(const V3& a) { return V3(-a.x, -a.y, -a.z); }
ATN_HD float dot(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
ATN_HD V3 cross(const V3& a, const V3& b)
{
    return V3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
ATN_HD float length(const V3& a) { return sqrtf(dot(a, a)); }
ATN_HD float squared_length(const V3& a) { return dot(a, a); }
ATN_HD V3 normalize(const V3& a) { return a * (1.0f / sqrtf(dot(a, a))); }
// std::max / std::min argument semantics (matters for NaN): max(a,b) = (a<b)?b:a, min(a,b) = (b<a)?b:a
ATN_HD float smax(float a, float b) { return (a < b) ? b : a; }
ATN_HD float smin(float a, float b) { return (b < a) ? b : a; }
ATN_HD float sclamp(float v, float lo, float hi) { return (v < lo) ? lo : ((hi < v) ? hi : v); }   // std::clamp
ATN_HD float saturate(float v) { return sclamp(v, 0.0f, 1.0f); }
ATN_HD float mixf(float x, float y, float a) { return x * (1 - a) + y * a; }                          // math.h:290-296
ATN_HD V3 mix3(const V3& x, const V3& y, float a) { return x * (1 - a) + y * a; }
ATN_HD float max3(const V3& v) { return smax(smax(v.x, v.y), v.z); }                                   // vec3.h:345-348
ATN_HD float min3(const V3& v) { return smin(smin(v.x, v.y), v.z); }
ATN_HD float luminance(const V3& c) { return 0.212639f * c.x + 0.71517f * c.y + 0.0721926f * c.z; }   // misc/color.h:61-73

ATN_HD int32_t f2i(float f) { int32_t i; memcpy(&i, &f, 4); return i; }
ATN_HD float i2f(int32_t i) { float f; memcpy(&f, &i, 4); return f; }

struct Ray { V3 org, dir; };

// aten::ray::Offset (src/libaten/math/ray.h:26-74)
ATN_HD_CALL V3 ray_offset(const V3& o, const V3& n)
{
    const float origin = 1.0f / 32.0f;
    const float float_scale = 1.0f / 65536.0f;
    const float int_scale = 256.0f;
    const int32_t ix = (int32_t)(int_scale * n.x);
    const int32_t iy = (int32_t)(int_scale * n.y);
    const int32_t iz = (int32_t)(int_scale * n.z);
    const V3 pi(i2f(f2i(o.x) + (o.x < 0.0f ? -ix : ix)),
                i2f(f2i(o.y) + (o.y < 0.0f ? -iy : iy)),
                i2f(f2i(o.z) + (o.z < 0.0f ? -iz : iz)));
    return V3(fabsf(o.x) < origin ? o.x + float_scale * n.x : pi.x,
              fabsf(o.y) < origin ? o.y + float_scale * n.y : pi.y,
              fabsf(o.z) < origin ? o.z + float_scale * n.z : pi.z);
}
// aten::ray(o, d) / aten::ray(o, d, normal) (ray.h:12-24)
ATN_HD Ray make_ray(const V3& o, const V3& d) { Ray r; r.dir = normalize(d); r.org = o; return r; }
ATN_HD Ray make_ray(const V3& o, const V3& d, const V3& n) { Ray r; r.dir = normalize(d); r.org = ray_offset(o, n); return r; }

// aten::mat4::apply / applyXYZ (src/libaten/math/mat4.h:171-221), row major
ATN_HD V3 mat_apply(const AtnMat4& m, const V3& p)
{
    return V3(m.m[0] * p.x + m.m[1] * p.y + m.m[2] * p.z + m.m[3],
              m.m[4] * p.x + m.m[5] * p.y + m.m[6] * p.z + m.m[7],
              m.m[8] * p.x + m.m[9] * p.y + m.m[10] * p.z + m.m[11]);
}
ATN_HD V3 mat_apply_w1(const AtnMat4& m, const V3& p)   // apply(vec4) with w = 1, xyz of the result
{
    return V3(m.m[0] * p.x + m.m[1] * p.y + m.m[2] * p.z + m.m[3] * 1.0f,
              m.m[4] * p.x + m.m[5] * p.y + m.m[6] * p.z + m.m[7] * 1.0f,
              m.m[8] * p.x + m.m[9] * p.y + m.m[10] * p.z + m.m[11] * 1.0f);
}
ATN_HD V3 mat_apply_xyz(const AtnMat4& m, const V3& p)
{
    return V3(m.m[0] * p.x + m.m[1] * p.y + m.m[2] * p.z,
              m.m[4] * p.x + m.m[5] * p.y + m.m[6] * p.z,
              m.m[8] * p.x + m.m[9] * p.y + m.m[10] * p.z);
}

// ---------------------------------------------------------------------------------------------
// A1: aten::CMJ (src/libaten/sampler/cmj.h:9-124) -- integer exact
// ---------------------------------------------------------------------------------------------
struct Sampler {
    uint32_t idx, dim, scramble;

    ATN_HD static uint32_t permute(uint32_t i, uint32_t l, uint32_t p)
    {
        uint32_t w = l - 1;
        w |= w >> 1; w |= w >> 2; w |= w >> 4; w |= w >> 8; w |= w >> 16;
        do {
            i ^= p; i *= 0xe170893d; i ^= p >> 16;
            i ^= (i & w) >> 4; i ^= p >> 8; i *= 0x0929eb3f; i ^= p >> 23;
            i ^= (i & w) >> 1; i *= 1 | p >> 27; i *= 0x6935fa69;
            i ^= (i & w) >> 11; i *= 0x74dcb303; i ^= (i & w) >> 2; i *= 0x9e501cc3;
            i ^= (i & w) >> 2; i *= 0xc860a3df; i &= w; i ^= i >> 5;
        } while (i >= l);
        return (i + p) % l;
    }
    ATN_HD static float randfloat(uint32_t i, uint32_t p)
    {
        i ^= p; i ^= i >> 17; i ^= i >> 10; i *= 0xb36534e5; i ^= i >> 12; i ^= i >> 21;
        i *= 0x93fc4795; i ^= 0xdf6e307f; i ^= i >> 17; i *= 1 | p >> 18;
        return i * (1.0f / 4294967808.0f);
    }
    // cmj(s, n, p).x with n = CMJ_DIM = 16 (cmj.h:100-111); nextSample() only uses .x
    ATN_HD_CALL float next()
    {
        const uint32_t ds = dim * scramble;
        const int32_t s = (int32_t)permute(idx, 256u, 0xa399d265u * ds);
        const int32_t n = 16;
        const int32_t p = (int32_t)ds;
        const int32_t sy = (int32_t)permute((uint32_t)(s / n), (uint32_t)n, (uint32_t)p * 0x63d83595u);
        const float jx = randfloat((uint32_t)s, (uint32_t)p * 0xa399d265u);
        dim++;
        return (s % n + (sy + jx) / n) / n;
    }
};

// A2: GeneratePath's sampler seeding (src/libaten/renderer/pathtracing/pathtracing_impl.h:75-87)
ATN_HD Sampler sampler_for_path(uint32_t frame_plus_sample, uint32_t rnd)
{
    Sampler s;
    const uint32_t scramble = rnd * 0x1fe3434fu * ((frame_plus_sample + 133u * rnd) / 256u);
    s.idx = frame_plus_sample % 256u;
    s.dim = 0;
    s.scramble = scramble;
    return s;
}

// A2: PinholeCamera::sample (src/libaten/camera/pinhole.cpp:97-118)
ATN_HD Ray camera_sample(const AtnCameraParameter& cam, float s, float t)
{
    s = 2.0f * s - 1.0f;
    t = 2.0f * t - 1.0f;
    V3 pos = s * V3(cam.u) + t * V3(cam.v);
    pos = pos + V3(cam.center);
    Ray r;
    r.dir = normalize(pos - V3(cam.origin));
    r.org = V3(cam.origin);
    return r;
}

// ---------------------------------------------------------------------------------------------
// A5: intersection primitives
// ---------------------------------------------------------------------------------------------
// intersectTriangle (src/libaten/math/intersect.h:45-89): Moeller-Trumbore without early outs.
struct TriHit { bool hit; float t, a, b; };
ATN_HD TriHit intersect_triangle(const V3& org, const V3& dir, const V3& v0, const V3& v1, const V3& v2)
{
    const V3 e1 = v1 - v0;
    const V3 e2 = v2 - v0;
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

// aabb::hit (src/libaten/math/aabb.h:61-86).  The +1e-6 bias on the direction is part of the
// reference's result and is reproduced; invdir / oxinvdir are per-ray constants.
struct RayInv { V3 invdir, oxinvdir; };
ATN_HD RayInv ray_inv(const V3& org, const V3& dir)
{
    RayInv ri;
    ri.invdir = V3(1.0f) / (dir + V3(1e-6f));
    ri.oxinvdir = (-org) * ri.invdir;
    return ri;
}
ATN_HD bool aabb_hit(const RayInv& ri, const V3& bmin, const V3& bmax, float t_min, float t_max, float* t0_out)
{
    const V3 f = bmax * ri.invdir + ri.oxinvdir;
    const V3 n = bmin * ri.invdir + ri.oxinvdir;
    const V3 tmax(smax(f.x, n.x), smax(f.y, n.y), smax(f.z, n.z));
    const V3 tmin(smin(f.x, n.x), smin(f.y, n.y), smin(f.z, n.z));
    const float t1 = smin(min3(tmax), t_max);
    const float t0 = smax(max3(tmin), t_min);
    *t0_out = t0;
    return t0 <= t1;
}

// sphere::hit (src/libaten/geometry/sphere.cpp:30-91); only reached from area-light sampling (F5)
ATN_HD bool is_close_ulps(float A, float B, int32_t maxUlps)       // math.h:306-340
{
    int32_t a = f2i(A); if (a < 0) a = (int32_t)(0x80000000u - (uint32_t)a);
    int32_t b = f2i(B); if (b < 0) b = (int32_t)(0x80000000u - (uint32_t)b);
    int32_t d = a - b; if (d < 0) d = -d;
    return d <= maxUlps;
}
ATN_HD bool sphere_hit(const V3& center, float radius, const Ray& r, float* t_out)
{
    const V3 p_o = center - r.org;
    const float b = dot(p_o, r.dir);
    const float D4 = b * b - dot(p_o, p_o) + radius * radius;
    if (D4 < 0.0f) return false;
    const float sq = sqrtf(D4);
    const float t1 = b - sq;
    const float t2 = b + sq;
    const bool close = is_close_ulps(fabsf(b), sq, 2500);
    if (t1 > kEps && !close) { *t_out = t1; return true; }
    if (t2 > kEps && !close) { *t_out = t2; return true; }
    return false;
}

// ---------------------------------------------------------------------------------------------
// Scene view: the flat arrays of AtnSceneBlob as the math sees them
// ---------------------------------------------------------------------------------------------
struct TexView { const AtnVec4* texels; int32_t w, h; int32_t filter, address; };   // aten::TextureFilterMode / TextureAddressMode

struct SceneView {
    const AtnObjectParameter* objects;
    const AtnMat4* matrices;
    const AtnMaterialParameter* materials;
    const AtnLightParameter* lights;
    const AtnTriangleParameter* triangles;
    const AtnVec4* positions;
    const AtnVec4* normals;
    const TexView* textures;
    int32_t light_count;
    int32_t material_count;
    AtnSceneConfig cfg;
    float scene_radius;      // ComputeDistanceToCoverBoundingSphere(30 deg) or 10000 (ibl.h:105-110)
    int32_t device_texture_fetch;   // 0: aten::texture::at (host rule, parity runs); 1: the GPU renderer's tex2DLod semantics per texture
    int32_t alpha_blending;  // SceneRenderingConfig::enable_alpha_blending (scene_rendering_config.h:21): shadow rays pass alpha surfaces    // Two float4 per light: conservative world-space bounds of an area light's object in the top layer (min, max; min.w != 0
    // when valid).  A shadow ray towards that light that hits anything clearly in front of these bounds is hidden whatever
    // else it would hit (the closest hit cannot be the light's object).  NULL in the CPU restatement.
    const AtnVec4* light_bounds;
};

ATN_HD AtnVec4 ld4(const AtnVec4* p)
{
#if defined(__CUDA_ARCH__)
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    AtnVec4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w; return r;
#else
    return *p;
#endif
}

// aten::texture::at (src/libaten/image/texture.cpp:64-75) + NormalizeToWrapRepeat (texture.h:197-208):
// the host (oracle) rule.  A8 / SURVEY Q12.
ATN_HD int32_t wrap_repeat(int32_t value, int32_t wrap)
{
    if (value > wrap) { const int32_t n = value / wrap; value -= n * wrap; }
    else if (value < 0) { int32_t n = value / wrap; if (n < 0) n = -n; value += (n + 1) * wrap; }
    return value;
}
// The fetch the reference's GPU renderer does instead (device sampleTexture, sample_texture.h:18-35: tex2DLod<float4> on a
// texture object with normalised coordinates and the texture's own filter and address mode, cudaTextureResource.cpp:210-236),
// written out in software after the texture-fetch formulas of the CUDA programming guide: with N texels along an axis and
// x = N * coordinate, point filtering reads T[floor(x)]; linear filtering reads T[i], T[i+1] with i = floor(x - 0.5) and the
// weight frac(x - 0.5) held in 8 fractional bits; wrap takes the fractional part of the coordinate and wraps indices, clamp
// clamps indices to the edge, mirror reflects every other period, border reads zeros outside.  PARITY UNPINNED: there is no
// GPU build of the reference on this box to compare with (no graphics stack for libidaten); pinned only to an independent
// restatement of the same published formulas (tests/test_texture_fetch.py).  Only mip level 0 exists (lod is ignored).
ATN_HD float tex_coord(float c, int32_t address)
{
    if (address == 0) return c - floorf(c);                                    // Wrap
    if (address == 2) { const float f = floorf(c); const float r = c - f; return ((int32_t)f & 1) ? 1.0f - r : r; }     // Mirror
    return c;                                                                   // Clamp, Border: as is
}
ATN_HD AtnVec4 tex_texel(const TexView& t, int32_t i, int32_t j)
{
    if (t.address == 0) {                                                       // Wrap: indices modulo the size
        i %= t.w; if (i < 0) i += t.w;
        j %= t.h; if (j < 0) j += t.h;
    }
    else if (t.address == 3) {                                                  // Border
        if (i < 0 || i >= t.w || j < 0 || j >= t.h) { AtnVec4 z; z.x = z.y = z.z = z.w = 0.0f; return z; }
    }
    else {                                                                      // Clamp, Mirror: to the edge
        i = i < 0 ? 0 : (i >= t.w ? t.w - 1 : i);
        j = j < 0 ? 0 : (j >= t.h ? t.h - 1 : j);
    }
    return ld4(&t.texels[(uint32_t)(j * t.w + i)]);
}
ATN_HD_CALL AtnVec4 tex_fetch_device(const TexView& t, float u, float v)
{
    const float x = tex_coord(u, t.address) * (float)t.w;
    const float y = tex_coord(v, t.address) * (float)t.h;
    if (t.filter != 1) return tex_texel(t, (int32_t)floorf(x), (int32_t)floorf(y));          // Point
    const float xb = x - 0.5f, yb = y - 0.5f;
    const float fi = floorf(xb), fj = floorf(yb);
    const float a = floorf((xb - fi) * 256.0f + 0.5f) * (1.0f / 256.0f);        // 1.8 fixed point
    const float b = floorf((yb - fj) * 256.0f + 0.5f) * (1.0f / 256.0f);
    const int32_t i = (int32_t)fi, j = (int32_t)fj;
    const AtnVec4 t00 = tex_texel(t, i, j), t10 = tex_texel(t, i + 1, j), t01 = tex_texel(t, i, j + 1), t11 = tex_texel(t, i + 1, j + 1);
    AtnVec4 r;
    r.x = (1 - a) * (1 - b) * t00.x + a * (1 - b) * t10.x + (1 - a) * b * t01.x + a * b * t11.x;
    r.y = (1 - a) * (1 - b) * t00.y + a * (1 - b) * t10.y + (1 - a) * b * t01.y + a * b * t11.y;
    r.z = (1 - a) * (1 - b) * t00.z + a * (1 - b) * t10.z + (1 - a) * b * t01.z + a * b * t11.z;
    r.w = (1 - a) * (1 - b) * t00.w + a * (1 - b) * t10.w + (1 - a) * b * t01.w + a * b * t11.w;
    return r;
}

ATN_HD_CALL AtnVec4 sample_texture(const SceneView& sc, int32_t texid, float u, float v, const AtnVec4& def)
{
    if (texid < 0) return def;
    const TexView t = sc.textures[texid];
    if (!t.texels) return def;
    if (sc.device_texture_fetch) return tex_fetch_device(t, u, v);
    const int32_t iu = (int32_t)(u * (t.w - 1));
    const int32_t iv = (int32_t)(v * (t.h - 1));
    const int32_t x = wrap_repeat(iu, t.w - 1);
    const int32_t y = wrap_repeat(iv, t.h - 1);
    return ld4(&t.texels[(uint32_t)(y * t.w + x)]);
}
ATN_HD AtnVec4 v4(float a) { AtnVec4 r; r.x = r.y = r.z = r.w = a; return r; }

// GetOrthoVector / GetTangentCoordinate (src/libaten/math/vec3.h:290-337)
ATN_HD_CALL void tangent_frame(const V3& n, V3* t_out, V3* b_out)
{
    V3 p;
    if (fabsf(n.z) > 0.0f) {
        const float k = sqrtf(n.y * n.y + n.z * n.z);
        p.x = 0; p.y = -n.z / k; p.z = n.y / k;
    }
    else {
        const float k = sqrtf(n.x * n.x + n.y * n.y);
        p.x = n.y / k; p.y = -n.x / k; p.z = 0;
    }
    p = normalize(p);
    const V3 b = cross(n, p);
    *t_out = cross(b, n);
    *b_out = b;
}

// applyNormalMap (src/libaten/material/sample_texture.h:64-87)
ATN_HD_CALL V3 apply_normal_map(const SceneView& sc, int32_t normal_map, const V3& org_nml, float u, float v)
{
    if (normal_map >= 0) {
        V3 nml = V3(sample_texture(sc, normal_map, u, v, v4(0.0f)));
        nml = 2.0f * nml - V3(1.0f);
        nml = normalize(nml);
        const V3 n = normalize(org_nml);
        V3 t, b;
        tangent_frame(n, &t, &b);
        V3 r = nml.z * n + nml.x * t + nml.y * b;
        return normalize(r);
    }
    return normalize(org_nml);
}

// ---------------------------------------------------------------------------------------------
// A7: hit evaluation
// ---------------------------------------------------------------------------------------------
struct HitRecord { V3 p; float area; V3 normal; int32_t mtrlid; float u, v; };

// triangle::EvaluateHitResult (src/libaten/geometry/triangle.h:70-120) +
// PolygonObject::evaluate_hit_result (PolygonObject.h:38-72) + evaluate_hit_result
// (EvaluateHitResult.h:11-72).  Sphere objects: sphere::EvaluateHitResult (sphere.cpp:93-109).
ATN_HD_CALL void evaluate_hit(const SceneView& sc, const AtnObjectParameter& obj, const Ray& ray,
                         int32_t primid, float a, float b, float t, int32_t mtrlid, HitRecord* rec)
{
    const bool inst = (obj.type == 1);
    const AtnObjectParameter& real = inst ? sc.objects[obj.object_id] : obj;
    const int32_t mtx_id = inst ? obj.mtx_id : -1;
    AtnMat4 I;
    for (int i = 0; i < 16; i++) I.m[i] = (i % 5 == 0) ? 1.0f : 0.0f;
    const AtnMat4 L2W = (mtx_id >= 0) ? sc.matrices[mtx_id] : I;

    if (real.type == 0) {
        const AtnTriangleParameter tri = sc.triangles[primid];
        const AtnVec4 p0 = ld4(&sc.positions[tri.idx[0]]);
        const AtnVec4 p1 = ld4(&sc.positions[tri.idx[1]]);
        const AtnVec4 p2 = ld4(&sc.positions[tri.idx[2]]);
        const AtnVec4 n0 = ld4(&sc.normals[tri.idx[0]]);
        const AtnVec4 n1 = ld4(&sc.normals[tri.idx[1]]);
        const AtnVec4 n2 = ld4(&sc.normals[tri.idx[2]]);
        const float c = 1 - a - b;
        rec->p = c * V3(p0) + a * V3(p1) + b * V3(p2);
        rec->normal = c * V3(n0) + a * V3(n1) + b * V3(n2);
        rec->u = c * p0.w + a * p1.w + b * p2.w;
        rec->v = c * n0.w + a * n1.w + b * n2.w;
        if (tri.needNormal > 0) {
            const V3 e01 = V3(p1) - V3(p0);
            const V3 e02 = V3(p2) - V3(p0);
            rec->normal = normalize(cross(e01, e02));
        }
        const float original_len = length(V3(p1) - V3(p0));
        const V3 q0 = mat_apply_w1(L2W, V3(p0));
        const V3 q1 = mat_apply_w1(L2W, V3(p1));
        const float scaled_len = length(q1 - q0);
        float ratio = scaled_len / original_len;
        ratio = ratio * ratio;
        rec->area = real.area * ratio;
    }
    else if (real.type == 2) {
        const V3 center(real.sphere_center[0], real.sphere_center[1], real.sphere_center[2]);
        rec->p = ray.org + t * ray.dir;
        rec->normal = (rec->p - center) / real.sphere_radius;
        const float radius = real.sphere_radius;
        rec->area = 4 * kPi * radius * radius;
        rec->u = 0; rec->v = 0;   // getUV is not needed by the callers on this path (light sampling)
    }
    rec->p = mat_apply(L2W, rec->p);
    rec->normal = normalize(mat_apply_xyz(L2W, rec->normal));
    rec->mtrlid = mtrlid;
}

// ---------------------------------------------------------------------------------------------
// A8/A9: materials
// ---------------------------------------------------------------------------------------------
struct Mtrl {          // the fields of aten::MaterialParameter (material.h:242-325) the path reads
    AtnVec4 baseColor;
    int32_t type;
    uint32_t attrib;
    int32_t id;
    int32_t isIdealRefraction;
    int32_t albedoMap, normalMap, roughnessMap;
    AtnStandardMaterialParameter s;
    ATN_HD bool is_emissive() const { return attrib & 1u; }
    ATN_HD bool is_singular() const { return attrib & 2u; }
    ATN_HD bool is_translucent() const { return attrib & 4u; }
};

// FillMaterial (src/libaten/material/material_impl.h:232-262)
ATN_HD Mtrl fill_material(const SceneView& sc, int32_t mtrl_id)
{
    Mtrl m;
    if (mtrl_id >= 0) {
        const AtnMaterialParameter& p = sc.materials[mtrl_id];
        m.baseColor = p.baseColor; m.type = p.type; m.attrib = p.attrib; m.id = p.id;
        m.isIdealRefraction = p.isIdealRefraction;
        m.albedoMap = p.albedoMap; m.normalMap = p.normalMap; m.roughnessMap = p.roughnessMap;
        m.s = p.standard;
    }
    else {
        // MaterialParameter(Diffuse, MaterialAttributeDiffuse), baseColor = 1
        m.baseColor.x = m.baseColor.y = m.baseColor.z = 1.0f; m.baseColor.w = 1.0f;
        m.type = ATN_MTRL_DIFFUSE; m.attrib = 0; m.id = 0; m.isIdealRefraction = 0;
        m.albedoMap = m.normalMap = m.roughnessMap = -1;
        m.s.ior = 1.0f; m.s.roughness = 0.5f; m.s.shininess = 1.0f; m.s.subsurface = 0.5f; m.s.metallic = 0.5f;
        m.s.specular = 0.5f; m.s.specularTint = 0.5f; m.s.anisotropic = 0.5f; m.s.sheen = 0.5f; m.s.sheenTint = 0.5f;
        m.s.clearcoat = 0.5f; m.s.clearcoatGloss = 0.5f;
    }
    return m;
}

struct MtrlSample { V3 dir; V3 bsdf; float pdf; };

// material::ComputeReflectVector (material.h:520-527)
ATN_HD V3 reflect_vec(const V3& wi, const V3& n) { V3 wo = wi - 2 * dot(wi, n) * n; return normalize(wo); }
// material::ComputeRefractVector (material.h:537-565)
ATN_HD V3 refract_vec(float ni, float nt, const V3& wi, const V3& n)
{
    const V3 w = -wi;
    V3 N = n;
    float costheta = dot(w, n);
    if (costheta < 0.0f) { const float tmp = ni; ni = nt; nt = tmp; costheta = -costheta; N = -N; }
    const float sintheta_2 = 1.0f - costheta * costheta;
    const float ni_nt = ni / nt;
    const float ni_nt_2 = ni_nt * ni_nt;
    V3 wo = (ni_nt * costheta - sqrtf(1.0f - ni_nt_2 * sintheta_2)) * N - ni_nt * w;
    return normalize(wo);
}
// material::ComputeSchlickFresnel (material.h:470-503, 505-514)
ATN_HD float schlick_f0_cos(float f0, float costheta)
{
    const float c = saturate(1 - costheta);
    const float c5 = c * c * c * c * c;
    return f0 + (1.0f - f0) * c5;
}
ATN_HD V3 schlick_f0_cos3(const V3& f0, float costheta)
{
    const float c = saturate(1 - costheta);
    const float c5 = c * c * c * c * c;
    return f0 + (V3(1.0f) - f0) * c5;
}
ATN_HD float schlick_fresnel(float ni, float nt, const V3& w, const V3& n)
{
    float costheta = dot(w, n);
    if (costheta < 0) { const float tmp = ni; ni = nt; nt = tmp; costheta = -costheta; }
    float f0 = (ni - nt) / (ni + nt);
    f0 = f0 * f0;
    return schlick_f0_cos(f0, costheta);
}

// Diffuse (src/libaten/material/diffuse.h:86-136)
ATN_HD float lambert_pdf(const V3& n, const V3& wo) { const float c = fabsf(dot(n, wo)); return c / kPi; }
ATN_HD_CALL V3 lambert_dir(const V3& n, float r1, float r2)
{
    const float costheta = sqrtf(1 - r1);
    const float sintheta = sqrtf(r1);
    const float phi = kPi2 * r2;
    const float cosphi = cosf(phi);
    const float sinphi = sinf(phi);
    V3 t, b;
    tangent_frame(n, &t, &b);
    V3 dir = t * sintheta * cosphi + b * sintheta * sinphi + n * costheta;
    return normalize(dir);
}
ATN_HD V3 lambert_brdf() { return V3(1.0f) / kPi; }

// MicrofacetGGX (src/libaten/material/ggx.cpp:101-274)
ATN_HD float ggx_D(const V3& m, const V3& n, float roughness)
{
    const float a = roughness;
    const float a2 = a * a;
    const float costheta = fabsf(dot(m, n));
    const float cos2 = costheta * costheta;
    const float denom = (a2 - 1) * cos2 + 1.0f;
    const float denom2 = denom * denom;
    return denom > 0 ? a2 / (kPi * denom2) : 0;
}
ATN_HD float ggx_lambda(float roughness, const V3& w, const V3& n)
{
    const float alpha = roughness;
    const float cos_theta = fabsf(dot(w, n));
    const float cos2 = cos_theta * cos_theta;
    const float sin2 = 1.0f - cos2;
    const float tan2 = sin2 / cos2;
    const float a2 = 1.0f / (alpha * alpha * tan2);
    return (-1.0f + sqrtf(1.0f + 1.0f / a2)) / 2.0f;
}
ATN_HD float ggx_G2(float roughness, const V3& view, const V3& light, const V3& n)
{
    const float li = ggx_lambda(roughness, view, n);
    const float lo = ggx_lambda(roughness, light, n);
    return 1.0f / (1.0f + li + lo);
}
ATN_HD float ggx_pdf_h(float roughness, const V3& n, const V3& m, const V3& wo)
{
    const float D = ggx_D(m, n, roughness);
    const float costheta = fabsf(dot(m, n));
    const float denom = 4 * fabsf(dot(wo, m));
    return denom > 0 ? (D * costheta) / denom : 0;
}
ATN_HD float ggx_pdf(float roughness, const V3& n, const V3& wi, const V3& wo)
{
    const V3 wh = normalize(-wi + wo);
    return ggx_pdf_h(roughness, n, wh, wo);
}
ATN_HD_CALL V3 ggx_sample_m(float roughness, const V3& n, float r1, float r2)
{
    const float a = roughness;
    float theta = atanf(a * sqrtf(r1 / (1 - r1)));
    theta = ((theta >= 0) ? theta : (theta + 2 * kPi));
    const float phi = 2 * kPi * r2;
    const float costheta = cosf(theta);
    const float sintheta = sinf(theta);
    const float cosphi = cosf(phi);
    const float sinphi = sinf(phi);
    V3 t, b;
    tangent_frame(n, &t, &b);
    V3 m = t * sintheta * cosphi + b * sintheta * sinphi + n * costheta;
    return normalize(m);
}
ATN_HD V3 ggx_sample_dir(float r1, float r2, float roughness, const V3& wi, const V3& n)
{
    const V3 m = ggx_sample_m(roughness, n, r1, r2);
    return reflect_vec(wi, m);
}
ATN_HD_CALL V3 ggx_brdf(float roughness, float ior, const V3& n, const V3& wi, const V3& wo)
{
    const V3 V = -wi;
    const V3 L = wo;
    const V3 N = n;
    const V3 H = normalize(L + V);
    const float NL = fabsf(dot(N, L));
    const float NV = fabsf(dot(N, V));
    const float D = ggx_D(H, N, roughness);
    const float G = ggx_G2(roughness, V, L, N);
    const float F = schlick_fresnel(1.0f, ior, L, H);
    const float denom = 4 * NL * NV;
    const float bsdf = denom > kEps ? F * G * D / denom : 0.0f;
    return V3(bsdf);
}

// MicrofacetBeckman (src/libaten/material/beckman.cpp:103-255)
ATN_HD float beckman_D(const V3& m, const V3& n, float roughness)
{
    const float costheta = fabsf(dot(m, n));
    if (costheta <= 0) return 0;
    const float cos2 = costheta * costheta;
    const float cos4 = cos2 * cos2;
    const float sintheta = sqrtf(1 - cos2);
    const float tantheta = sintheta / costheta;
    const float tan2 = tantheta * tantheta;
    const float a2 = roughness * roughness;
    float D = 1.0f / (kPi * a2 * cos4);
    D *= expf(-tan2 / a2);
    return D;
}
ATN_HD_COLD float beckman_pdf(float roughness, const V3& n, const V3& wi, const V3& wo)
{
    const V3 wh = normalize(-wi + wo);
    const float costheta = fabsf(dot(wh, n));
    const float D = beckman_D(wh, n, roughness);
    const float denom = 4 * fabsf(dot(wo, wh));
    return denom > 0 ? (D * costheta) / denom : 0;
}
ATN_HD_COLD V3 beckman_sample_dir(float roughness, const V3& wi, const V3& n, float r1, float r2)
{
    const float a2 = roughness * roughness;
    const float theta = atanf(sqrtf(-a2 * logf(1.0f - r1 * 0.99f)));
    const float phi = kPi2 * r2;
    const float costheta = cosf(theta);
    const float sintheta = sinf(theta);
    const float cosphi = cosf(phi);
    const float sinphi = sinf(phi);
    V3 t, b;
    tangent_frame(n, &t, &b);
    V3 m = t * sintheta * cosphi + b * sintheta * sinphi + n * costheta;
    m = normalize(m);
    return reflect_vec(wi, m);
}
ATN_HD float beckman_G1(float roughness, const V3& v, const V3& n)
{
    const float costheta = saturate(fabsf(dot(v, n)));
    const float sintheta = sqrtf(1.0f - costheta * costheta);
    const float tantheta = sintheta / costheta;
    const float a = 1.0f / (roughness * tantheta);
    const float a2 = a * a;
    if (a < 1.6f) return (3.535f * a + 2.181f * a2) / (1.0f + 2.276f * a + 2.577f * a2);
    return 1.0f;
}
ATN_HD_COLD V3 beckman_brdf(float roughness, float ior, const V3& n, const V3& wi, const V3& wo)
{
    const V3 V = -wi;
    const V3 L = wo;
    const V3 N = n;
    const V3 H = normalize(L + V);
    const float NL = fabsf(dot(N, L));
    const float NV = fabsf(dot(N, V));
    const float D = beckman_D(H, N, roughness);
    const float G2 = beckman_G1(roughness, V, N) * beckman_G1(roughness, L, N);
    const float F = schlick_fresnel(1.0f, ior, L, H);
    const float denom = 4 * NL * NV;
    const float bsdf = denom > kEps ? F * G2 * D / denom : 0.0f;
    return V3(bsdf);
}

// specular (src/libaten/material/specular.h:60-103)
ATN_HD V3 specular_brdf(const V3& wo, const V3& n)
{
    const float c = dot(n, wo);
    const float bsdf = c == 0.0f ? 0.0f : 1.0f / c;
    return V3(bsdf);
}

// refraction::SampleRefraction / ComputeBRDF (src/libaten/material/refraction.cpp:63-161)
ATN_HD V3 refraction_brdf(float ni, float nt, const V3& wo, const V3& n, float T)
{
    const float c = fabsf(dot(wo, n));
    const float nt_ni = nt / ni;
    const float bsdf = c == 0.0f ? 0.0f : (nt_ni * nt_ni) * T / c;
    return V3(bsdf);
}
ATN_HD_COLD void refraction_sample(MtrlSample* result, Sampler* sampler, const Mtrl& param, const V3& n, const V3& wi)
{
    float ni = 1.0f;
    float nt = param.s.ior;
    const V3 V = -wi;
    V3 N = n;
    const bool is_enter = dot(V, N) >= 0.0f;
    if (!is_enter) { N = -n; const float tmp = ni; ni = nt; nt = tmp; }
    const float ni_nt = ni / nt;
    const float cos_i = dot(V, N);
    const float cos_t_2 = 1.0f - (ni_nt * ni_nt * (1.0f - cos_i * cos_i));
    if (cos_t_2 < 0.0f) { const float tmp = ni; ni = nt; nt = tmp; N = -N; }
    V3 wo = refract_vec(ni, nt, wi, N);
    const float F = schlick_fresnel(ni, nt, wo, N);
    const float R = F;
    const float T = 1 - R;
    if (param.isIdealRefraction) {
        result->pdf = 1.0f; result->dir = wo; result->bsdf = refraction_brdf(ni, nt, wo, N, T);
        return;
    }
    const float prob = 0.25f + 0.5f * R;
    const float u = sampler->next();
    if (u < prob) {
        wo = reflect_vec(wi, N);
        const float c = fabsf(dot(wo, N));
        const float bsdf = c == 0.0f ? 0.0f : R / c;
        result->pdf = prob; result->dir = wo; result->bsdf = V3(bsdf);
    }
    else {
        result->pdf = 1.0f - prob; result->dir = wo; result->bsdf = refraction_brdf(ni, nt, wo, N, T);
    }
}

// DisneyBRDF (src/libaten/material/disney_brdf.cpp:40-555); Q9 (clearcoat pdf evaluated with
// `roughness` in bsdf() and in sample()) is reproduced.
ATN_HD V3 disney_ctint(const V3& base)
{
    const float Y = dot(base, V3(0.3f, 0.6f, 0.1f));
    return Y > 0 ? base / Y : V3(1.0f);
}
ATN_HD float disney_schlick(float u)
{
    const float m = sclamp(1.0f - u, 0.0f, 1.0f);
    const float m2 = m * m;
    return m2 * m2 * m;
}
ATN_HD V3 disney_diffuse_brdf(const V3& base, float roughness, float subsurface, const V3& V, const V3& L, const V3& N)
{
    const V3 H = normalize(V + L);
    const float LdotH = dot(L, H);
    const float NdotV = dot(V, N);
    const float NdotL = dot(L, N);
    const float FV = disney_schlick(dot(V, N));
    const float FL = disney_schlick(dot(L, N));
    const float Fd90 = 0.5f + 2 * LdotH * LdotH * roughness;
    float fd = mixf(1.0f, Fd90, FL) * mixf(1.0f, Fd90, FV);
    const float Fss90 = LdotH * LdotH * roughness;
    const float Fss = mixf(1.0f, Fss90, FL) * mixf(1.0f, Fss90, FV);
    const float ss = 1.25f * (Fss * (1.0f / (NdotL + NdotV) - 0.5f) + 0.5f);
    fd = mixf(fd, ss, subsurface);
    return base / kPi * fd;
}
ATN_HD V3 disney_sheen_brdf(const V3& base, float sheen, float sheen_tint, const V3& V, const V3& L)
{
    const V3 H = normalize(V + L);
    const V3 Ctint = disney_ctint(base);
    const V3 Csheen = mix3(V3(1.0f), Ctint, sheen_tint);
    const float FH = disney_schlick(dot(L, H));
    return sheen * Csheen * FH;
}
ATN_HD float disney_gtr1(float roughness, float NdotH)
{
    const float a = roughness;
    if (a >= 1) return 1 / kPi;
    const float a2 = a * a;
    const float t = 1.0f + (a2 - 1.0f) * NdotH * NdotH;
    return (a2 - 1) / (kPi * logf(a2) * t);
}
ATN_HD V3 disney_clearcoat_brdf(float clearcoat, const V3& V, const V3& L)
{
    const V3 H = normalize(V + L);
    const float LdotH = dot(L, H);
    const float FH = disney_schlick(fabsf(LdotH));
    const float F = mixf(0.04f, 1.0f, FH);
    return V3(0.25f * clearcoat * F);
}
ATN_HD float disney_clearcoat_pdf(float clearcoat_gloss, const V3& V, const V3& L, const V3& N)
{
    const V3 H = normalize(V + L);
    const float NdotH = dot(N, H);
    const float a = mixf(0.1f, 0.001f, clearcoat_gloss);
    const float D = disney_gtr1(a, NdotH);
    const float costheta = fabsf(dot(H, N));
    const float denom = 4 * fabsf(dot(L, H));
    return denom > 0 ? (D * costheta) / denom : 0;
}
ATN_HD V3 disney_specular_brdf(const V3& base, float roughness, float metallic, float specular, float specular_tint,
                               const V3& V, const V3& L, const V3& N)
{
    const V3 H = normalize(V + L);
    const V3 Ctint = disney_ctint(base);
    const V3 Cspec = mix3(V3(1.0f), Ctint, specular_tint);
    const V3 F_s0 = mix3(0.08f * specular * Cspec, base, metallic);
    const V3 F = mix3(F_s0, V3(1.0f), dot(L, H));
    const float D = ggx_D(H, N, roughness);
    const float G = ggx_G2(roughness, V, L, N);
    const float NdotL = fabsf(dot(N, L));
    const float NdotV = fabsf(dot(N, V));
    const float denom = 4 * NdotV * NdotL;
    return denom > 0 ? F * D * G / denom : V3(0.0f);
}
ATN_HD float disney_specular_pdf(float roughness, const V3& V, const V3& L, const V3& N)
{
    const V3 H = normalize(V + L);
    return ggx_pdf_h(roughness, N, H, L);
}
ATN_HD void disney_weights(float w[4], const V3& base, float metallic, float sheen, float specular, float clearcoat)
{
    const float lum = luminance(base);
    w[0] = lum * (1 - metallic);
    w[1] = sheen * (1 - metallic);
    w[2] = mixf(specular, 1.0f, metallic);
    w[3] = 0.25f * clearcoat;
    float norm = 0.0f;
    for (int i = 0; i < 4; i++) norm += w[i];
    if (norm > 0) for (int i = 0; i < 4; i++) w[i] /= norm;
}
// DisneyBRDF::bsdf (disney_brdf.cpp:371-436) and the shared tail of ::sample (:520-553)
ATN_HD void disney_eval_rest(const Mtrl& m, float w[4], const V3& V, const V3& wo, const V3& N,
                             V3* d, V3* sh, V3* sp, V3* cc, float* pdf)
{
    const V3 base(m.baseColor);
    if (w[0] > 0.0f) { *d = disney_diffuse_brdf(base, m.s.roughness, m.s.subsurface, V, wo, N); *pdf += w[0] * lambert_pdf(N, wo); }
    if (w[1] > 0.0f) { *sh = disney_sheen_brdf(base, m.s.sheen, m.s.sheenTint, V, wo); *pdf += w[1] * (1 / kPi); }
    if (w[2] > 0.0f) { *sp = disney_specular_brdf(base, m.s.roughness, m.s.metallic, m.s.specular, m.s.specularTint, V, wo, N);
                       *pdf += w[2] * disney_specular_pdf(m.s.roughness, V, wo, N); }
    if (w[3] > 0.0f) { *cc = disney_clearcoat_brdf(m.s.clearcoat, V, wo); *pdf += w[3] * disney_clearcoat_pdf(m.s.roughness, V, wo, N); }
}
ATN_HD_COLD float disney_pdf(const Mtrl& m, const V3& n, const V3& wi, const V3& wo)     // disney_brdf.cpp:338-369
{
    float w[4];
    disney_weights(w, V3(m.baseColor), m.s.metallic, m.s.sheen, m.s.specular, m.s.clearcoat);
    const V3 V = -wi;
    float pdf = 0.0f;
    pdf += w[0] * lambert_pdf(n, wo);
    pdf += w[1] * (1 / kPi);
    pdf += w[2] * disney_specular_pdf(m.s.roughness, V, wo, n);
    pdf += w[3] * disney_clearcoat_pdf(m.s.clearcoatGloss, V, wo, n);
    return pdf;
}
ATN_HD_COLD MtrlSample disney_bsdf(const Mtrl& m, const V3& n, const V3& wi, const V3& wo)
{
    float w[4];
    disney_weights(w, V3(m.baseColor), m.s.metallic, m.s.sheen, m.s.specular, m.s.clearcoat);
    const V3 V = -wi;
    float pdf = 0.0f;
    V3 d(0.0f), sh(0.0f), sp(0.0f), cc(0.0f);
    // bsdf() accumulates prob * weight (operands swapped w.r.t. sample(); multiplication commutes)
    disney_eval_rest(m, w, V, wo, n, &d, &sh, &sp, &cc, &pdf);
    MtrlSample r;
    r.bsdf = (1 - m.s.metallic) * (d + sh) + sp + cc;
    r.pdf = pdf;
    r.dir = V3(0.0f);
    return r;
}
ATN_HD_COLD void disney_sample(MtrlSample* result, const Mtrl& m, const V3& n, const V3& wi, Sampler* sampler)
{
    const float r1 = sampler->next();
    const float r2 = sampler->next();
    const float r3 = sampler->next();
    const V3 base(m.baseColor);
    float w[4];
    disney_weights(w, base, m.s.metallic, m.s.sheen, m.s.specular, m.s.clearcoat);
    float cdf[4];
    cdf[0] = w[0]; cdf[1] = cdf[0] + w[1]; cdf[2] = cdf[1] + w[2]; cdf[3] = cdf[2] + w[3];
    const V3 V = -wi;
    const V3 N = n;
    V3 wo;
    float pdf = 0;
    V3 d(0.0f), sh(0.0f), sp(0.0f), cc(0.0f);
    if (r3 < cdf[0]) {
        wo = lambert_dir(N, r1, r2);
        d = disney_diffuse_brdf(base, m.s.roughness, m.s.subsurface, V, wo, N);
        pdf = lambert_pdf(N, wo); pdf *= w[0]; w[0] = 0.0f;
    }
    else if (r3 < cdf[1]) {
        wo = lambert_dir(N, r1, r2);
        sh = disney_sheen_brdf(base, m.s.sheen, m.s.sheenTint, V, wo);
        pdf = 1 / kPi; pdf *= w[1]; w[1] = 0.0f;
    }
    else if (r3 < cdf[2]) {
        wo = ggx_sample_dir(r1, r2, m.s.roughness, -V, N);
        sp = disney_specular_brdf(base, m.s.roughness, m.s.metallic, m.s.specular, m.s.specularTint, V, wo, N);
        pdf = disney_specular_pdf(m.s.roughness, V, wo, N); pdf *= w[2]; w[2] = 0.0f;
    }
    else {
        const float a = mixf(0.1f, 0.001f, m.s.clearcoatGloss);
        const V3 mm = ggx_sample_m(a, N, r1, r2);
        wo = reflect_vec(-V, mm);
        cc = disney_clearcoat_brdf(m.s.clearcoat, V, wo);
        pdf = disney_clearcoat_pdf(m.s.roughness, V, wo, N); pdf *= w[3]; w[3] = 0.0f;
    }
    disney_eval_rest(m, w, V, wo, N, &d, &sh, &sp, &cc, &pdf);
    result->pdf = pdf;
    result->bsdf = (1 - m.s.metallic) * (d + sh) + sp + cc;
    result->dir = wo;
}

ATN_HD float roughness_of(const SceneView& sc, const Mtrl& m, float u, float v)
{
    return sample_texture(sc, m.roughnessMap, u, v, v4(m.s.roughness)).x;
}

// material::sampleMaterial (src/libaten/material/material_impl.h:24-87), host variant
ATN_HD void sample_material(MtrlSample* r, const SceneView& sc, const Mtrl& m, const V3& normal, const V3& wi,
                            Sampler* sampler, float u, float v)
{
    switch (m.type) {
    case ATN_MTRL_SPECULAR:
        r->dir = reflect_vec(wi, normal); r->pdf = 1.0f; r->bsdf = specular_brdf(r->dir, normal);
        break;
    case ATN_MTRL_REFRACTION:
        refraction_sample(r, sampler, m, normal, wi);
        break;
    case ATN_MTRL_GGX: {
        const float rough = roughness_of(sc, m, u, v);
        const float r1 = sampler->next();
        const float r2 = sampler->next();
        r->dir = ggx_sample_dir(r1, r2, rough, wi, normal);
        r->pdf = ggx_pdf(rough, normal, wi, r->dir);
        r->bsdf = ggx_brdf(rough, m.s.ior, normal, wi, r->dir);
        break; }
    case ATN_MTRL_BECKMAN: {
        const float rough = roughness_of(sc, m, u, v);
        const float r1 = sampler->next();
        const float r2 = sampler->next();
        r->dir = beckman_sample_dir(rough, wi, normal, r1, r2);
        r->pdf = beckman_pdf(rough, normal, wi, r->dir);
        r->bsdf = beckman_brdf(rough, m.s.ior, normal, wi, r->dir);
        break; }
    case ATN_MTRL_DISNEY:
        disney_sample(r, m, normal, wi, sampler);
        break;
    case ATN_MTRL_EMISSIVE:   // emissive::sample == Diffuse (emissive.h:30-83)
    case ATN_MTRL_DIFFUSE:
    default: {                // unsupported types fall back to Diffuse like the reference's default branch
        const float r1 = sampler->next();
        const float r2 = sampler->next();
        r->dir = lambert_dir(normal, r1, r2);
        r->pdf = lambert_pdf(normal, r->dir);
        r->bsdf = lambert_brdf();
        break; }
    }
}
// material::samplePDF (material_impl.h:89-144)
ATN_HD float material_pdf(const SceneView& sc, const Mtrl& m, const V3& normal, const V3& wi, const V3& wo, float u, float v)
{
    switch (m.type) {
    case ATN_MTRL_SPECULAR: return 1.0f;
    case ATN_MTRL_REFRACTION: return 1.0f;
    case ATN_MTRL_GGX: return ggx_pdf(roughness_of(sc, m, u, v), normal, wi, wo);
    case ATN_MTRL_BECKMAN: return beckman_pdf(roughness_of(sc, m, u, v), normal, wi, wo);
    case ATN_MTRL_DISNEY: return disney_pdf(m, normal, wi, wo);
    default: return lambert_pdf(normal, wo);
    }
}
// material::sampleBSDF (material_impl.h:146-206): returns bsdf and (Disney only) pdf
ATN_HD MtrlSample material_bsdf(const SceneView& sc, const Mtrl& m, const V3& normal, const V3& wi, const V3& wo, float u, float v)
{
    MtrlSample r;
    r.pdf = 0.0f; r.dir = V3(0.0f);
    switch (m.type) {
    case ATN_MTRL_SPECULAR: r.bsdf = specular_brdf(wo, normal); break;
    case ATN_MTRL_REFRACTION: r.bsdf = V3(0.0f); break;
    case ATN_MTRL_GGX: r.bsdf = ggx_brdf(roughness_of(sc, m, u, v), m.s.ior, normal, wi, wo); break;
    case ATN_MTRL_BECKMAN: r.bsdf = beckman_brdf(roughness_of(sc, m, u, v), m.s.ior, normal, wi, wo); break;
    case ATN_MTRL_DISNEY: r = disney_bsdf(m, normal, wi, wo); break;
    default: r.bsdf = lambert_brdf(); break;
    }
    return r;
}

// ---------------------------------------------------------------------------------------------
// A10: lights
// ---------------------------------------------------------------------------------------------
struct LightSample {
    V3 pos, dir; float dist; V3 nml; V3 color; float pdf; uint32_t attrib;
};

// Background::ConvertDirectionToUV (src/libaten/renderer/background.h:88-127)
ATN_HD void dir_to_uv(const V3& dir, float* u, float* v)
{
    const float temp = atan2f(dir.x, dir.z);
    const float r = length(dir);
    const float phi = (float)((temp >= 0) ? temp : (temp + 2 * kPi));
    const float theta = acosf(dir.y / r);
    *u = phi / (2 * kPi);
    *v = 1 - theta / kPi;
}
// Background::SampleFromRay (background.h:34-62)
ATN_HD_CALL AtnVec4 background_sample(const SceneView& sc, const V3& dir)
{
    if (sc.cfg.envmap_tex_idx < 0 || !sc.cfg.enable_env_map) {
        AtnVec4 c; c.x = sc.cfg.bg_color.x; c.y = sc.cfg.bg_color.y; c.z = sc.cfg.bg_color.z; c.w = 0.0f;
        return c;
    }
    float u, v;
    dir_to_uv(dir, &u, &v);
    AtnVec4 c = sample_texture(sc, sc.cfg.envmap_tex_idx, u, v, v4(1.0f));
    c.x = c.x * sc.cfg.multiplyer; c.y = c.y * sc.cfg.multiplyer; c.z = c.z * sc.cfg.multiplyer; c.w = c.w * sc.cfg.multiplyer;
    return c;
}
// ImageBasedLight::samplePdf (src/libaten/light/ibl.h:46-58)
ATN_HD float ibl_pdf(const V3& clr, float avg_illum)
{
    float pdf = luminance(clr) / avg_illum;
    pdf /= (2.0f * kPi);
    return pdf;
}

// triangle::SamplePosAndNormal (triangle.h:122-162) + PolygonObject::SamplePosAndNormal
// (PolygonObject.h:114-156) + AreaLight::sample (arealight.h:37-143)
ATN_HD_CALL bool area_light_sample(LightSample* result, const AtnLightParameter& light, const SceneView& sc,
                              const V3& org, Sampler* sampler)
{
    if (light.arealight_objid < 0) return false;
    const AtnObjectParameter& obj = sc.objects[light.arealight_objid];
    const bool inst = (obj.type == 1);
    const AtnObjectParameter& real = inst ? sc.objects[obj.object_id] : obj;
    HitRecord rec;
    if (real.type == 0) {
        const float r = sampler->next();
        uint32_t tri_idx = (uint32_t)(real.triangle_num * r);
        tri_idx += real.triangle_id;
        const AtnTriangleParameter tri = sc.triangles[tri_idx];
        const AtnVec4 p0 = ld4(&sc.positions[tri.idx[0]]);
        const AtnVec4 p1 = ld4(&sc.positions[tri.idx[1]]);
        const AtnVec4 p2 = ld4(&sc.positions[tri.idx[2]]);
        const float r0 = sampler->next();
        const float r1 = sampler->next();
        const float a = sqrtf(r0) * (1.0f - r1);
        const float b = sqrtf(r0) * r1;
        const V3 pos = (1 - a - b) * V3(p0) + a * V3(p1) + b * V3(p2);   // LOCAL space, as in the reference
        const V3 dir = pos - org;
        const Ray ray = make_ray(org, dir);
        evaluate_hit(sc, obj, ray, (int32_t)tri_idx, a, b, length(dir), -1, &rec);
    }
    else if (real.type == 2) {
        // sphere::SamplePosAndNormal (sphere.cpp:111-150) then sphere::hit
        const float r1 = sampler->next();
        const float r2 = sampler->next();
        const float rad = real.sphere_radius;
        const float z = 2.0f * r1 - 1.0f;
        const float sin_theta = sqrtf(1 - z * z);
        const float phi = 2 * kPi * r2;
        const float x = cosf(phi) * sin_theta;
        const float y = sinf(phi) * sin_theta;
        V3 d = normalize(V3(x, y, z));
        const V3 center(real.sphere_center[0], real.sphere_center[1], real.sphere_center[2]);
        const V3 pos = center + d * (rad + kEps);
        const Ray ray = make_ray(org, pos - org);
        float t;
        if (!sphere_hit(center, rad, ray, &t)) return false;
        evaluate_hit(sc, obj, ray, -1, 0, 0, t, -1, &rec);
    }
    else {
        return false;
    }
    // AreaLight::sample(hrec, ...) (arealight.h:42-58)
    result->pos = rec.p;
    result->pdf = 1 / rec.area;
    result->dir = rec.p - org;
    result->dist = length(result->dir);
    result->dir = normalize(result->dir);
    result->nml = rec.normal;
    const float lum = light.scale * light.intensity / rec.area;
    result->color = V3(light.light_color) * lum;
    return true;
}

// Light::sample dispatch (src/libaten/light/light_impl.h:13-44)
ATN_HD_CALL void light_sample(LightSample* r, const AtnLightParameter& light, const SceneView& sc,
                         const V3& org, const V3& nml, Sampler* sampler)
{
    r->pos = V3(0.0f); r->dir = V3(0.0f); r->dist = 0.0f; r->nml = V3(0.0f); r->color = V3(0.0f); r->pdf = 0.0f;
    switch (light.type) {
    case 0:
        area_light_sample(r, light, sc, org, sampler);
        break;
    case 1: {   // ImageBasedLight::sample (src/libaten/light/ibl.h:71-133)
        const float r1 = sampler->next();
        const float r2 = sampler->next();
        r->dir = lambert_dir(nml, r1, r2);
        float u, v;
        dir_to_uv(r->dir, &u, &v);
        r->pos = org + sc.scene_radius * r->dir;
        r->nml = -normalize(r->dir);
        r->pdf = 1.0f / (2.0f * kPi);
        r->dist = 1.0f;
        const AtnVec4 lum = sample_texture(sc, light.envmapidx, u, v, v4(1.0f));
        r->color = light.scale * V3(lum);
        break; }
    case 2: {   // DirectionalLight::sample (directionallight.h:41-62)
        // NOTE: the reference normalises the *vec4* param.dir (vec4.h:283-288), whose w is 1
        // (LightParameter's default), so dir/nml come out with length 1/sqrt(2).  Reproduced.
        r->pdf = 1.0f;
        const AtnVec4 d4 = light.dir;
        const float inv4 = 1.0f / sqrtf(d4.x * d4.x + d4.y * d4.y + d4.z * d4.z + d4.w * d4.w);
        const V3 n4(inv4 * d4.x, inv4 * d4.y, inv4 * d4.z);
        r->dir = -n4;
        r->nml = n4;
        r->pos = org + 100000.0f * 0.5f * r->dir;
        r->color = V3(light.light_color) * light.scale * light.intensity;
        r->dist = 1.0f;
        break; }
    case 3: {   // PointLight::sample (pointlight.h:41-60)
        r->pdf = 1.0f;
        r->dir = V3(light.pos) - org;
        r->dist = length(r->dir);
        r->dir = normalize(r->dir);
        r->pos = V3(light.pos);
        r->nml = normalize(-r->dir);
        const float dist2 = r->dist * r->dist;
        r->color = V3(light.light_color) * light.scale * light.intensity / dist2;
        break; }
    case 4: {   // SpotLight::sample (spotlight.h:58-91)
        r->pdf = 1.0f;
        r->pos = V3(light.pos);
        r->nml = V3(light.dir);
        r->dir = V3(light.pos) - org;
        r->dist = length(r->dir);
        r->dir = normalize(r->dir);
        const V3 dir_to_light = -r->dir;
        const float rho = dot(V3(light.dir), dir_to_light);
        const float cosHalfInner = cosf(light.innerAngle * 0.5f);
        const float cosHalfOuter = cosf(light.outerAngle * 0.5f);
        if (rho > cosHalfOuter) {
            float att = (rho - cosHalfOuter) / (cosHalfInner - cosHalfOuter);
            att = sclamp(att, 0.0f, 1.0f);
            const float dist2 = r->dist * r->dist;
            r->color = light.scale * V3(light.light_color) * att * light.intensity / dist2;
        }
        else {
            r->pdf = 0.0f;
            r->color = V3(0.0f);
        }
        break; }
    default:
        break;
    }
    r->attrib = light.attrib;
}

// ComputeRadianceNEE (src/libaten/renderer/pathtracing/pathtracing_nee_impl.h:23-95)
ATN_HD_CALL bool radiance_nee(V3* out, const SceneView& sc, const V3& wi, const V3& nml, const Mtrl& m,
                         float u, float v, float light_select_prob, const LightSample& ls)
{
    const float cosShadow = dot(nml, ls.dir);
    float path_pdf = material_pdf(sc, m, nml, wi, ls.dir, u, v);
    const MtrlSample ev = material_bsdf(sc, m, nml, wi, ls.dir, u, v);
    if (ev.pdf > 0) path_pdf = ev.pdf;
    const V3 bsdf = ev.bsdf;
    const V3 emit = ls.color;
    const float cosLight = dot(ls.nml, -ls.dir);
    const bool infinite = (ls.attrib & 2u) != 0;
    const bool singular = (ls.attrib & 1u) != 0;
    float dist2 = ls.dist * ls.dist;
    dist2 = (infinite || singular) ? 1.0f : dist2;
    if (cosShadow >= 0 && cosLight >= 0 && dist2 > 0 && path_pdf > 0.0f && ls.pdf > 0.0f) {
        if (!infinite) path_pdf = path_pdf * cosLight / dist2;
        const float misW = singular ? 1.0f : (ls.pdf * light_select_prob) / ((ls.pdf * light_select_prob) + path_pdf);
        const float G = infinite ? cosShadow * cosLight : cosShadow * cosLight / dist2;
        *out = (misW * bsdf * emit * G / ls.pdf) / light_select_prob;
        return true;
    }
    return false;
}

}  // namespace atn

// ---------------------------------------------------------------------------------------------
// A10/A11/A12: the per-hit shade step and the miss step, shared by the CUDA shade kernel and the
// CPU restatement.  State is passed in registers; the callers own the SoA arrays.
// ---------------------------------------------------------------------------------------------
namespace atn {

struct PathState {
    V3 throughput; float pdfb;
    V3 contrib;
    Sampler sampler;
    bool terminated;
    bool singular;
    bool last_specular = false;   // PathAttribute::last_hit_mtrl_idx names a Specular material (read by the SVGF shade only)
};

struct ShadowRayOut {
    bool active;
    bool sampled = false;   // a light was sampled: FillShadowRay reset ShadowRay::lightcontrib (read again by ReSTIR's visibility step)
    V3 org; float dist;
    V3 dir; int32_t light_id;
    V3 lightcontrib;
};

// AOVs of the first hit (A8/A12): FillBasicAOVs / FillBasicAOVsIfHitMiss
// (src/libaten/renderer/aov.h:159-199).  `w2c` is the world-to-clip matrix; the reference's PT kernel
// passes the identity (src/libidaten/kernel/pathtracing_impl.cu:102-111), which makes depth = 1.
ATN_HD void fill_basic_aovs(AtnVec4* normal_depth, const V3& normal, const V3& p, const AtnMat4& w2c,
                            AtnVec4* albedo_meshid, const AtnVec4& albedo, int32_t meshid)
{
    // pos = mtx_W2C.apply(vec4(p, 1)); only .w is used (mat4.h:171-181)
    const float w = w2c.m[12] * p.x + w2c.m[13] * p.y + w2c.m[14] * p.z + w2c.m[15] * 1.0f;
    normal_depth->x = normal.x; normal_depth->y = normal.y; normal_depth->z = normal.z; normal_depth->w = w;
    albedo_meshid->x = albedo.x; albedo_meshid->y = albedo.y; albedo_meshid->z = albedo.z;
    albedo_meshid->w = (float)meshid;
}
ATN_HD void fill_basic_aovs_miss(AtnVec4* normal_depth, AtnVec4* albedo_meshid, const AtnVec4& bg)
{
    normal_depth->x = 0.0f; normal_depth->y = 0.0f; normal_depth->z = 0.0f; normal_depth->w = -1.0f;
    albedo_meshid->x = bg.x; albedo_meshid->y = bg.y; albedo_meshid->z = bg.z; albedo_meshid->w = -1.0f;
}
// Camera::ComputePixelWidthAtDistance (src/libaten/camera/camera.h:182-197)
ATN_HD float pixel_width_at_distance(const AtnCameraParameter& cam, float distance)
{
    distance = fabsf(distance);
    float hfov = cam.vfov * cam.height / (float)(cam.width);
    hfov = kPi * hfov / 180.0f;
    const float half_width = tanf(hfov / 2) * distance;
    const float width = half_width * 2;
    return width / (float)(cam.width);
}

// ShadeMiss (src/libaten/renderer/pathtracing/pathtracing_impl.h:112-176)
ATN_HD void shade_miss(const SceneView& sc, const AtnCameraParameter& cam, int32_t ix, int32_t iy,
                       int32_t width, int32_t height, int32_t bounce, const Ray& ray, PathState* st,
                       AtnVec4* aov_normal_depth = nullptr, AtnVec4* aov_albedo_meshid = nullptr)
{
    V3 dir = ray.dir;
    if (bounce == 0) {
        const float s = ix / (float)(width);
        const float t = iy / (float)(height);
        dir = camera_sample(cam, s, t).dir;
    }
    const AtnVec4 emit4 = background_sample(sc, dir);
    const V3 emit(emit4);
    float misW = 1.0f;
    if (bounce == 0 || (bounce == 1 && st->singular)) {
        // Export bg color to the albedo buffer (pathtracing_impl.h:150-158)
        if (aov_normal_depth && aov_albedo_meshid) fill_basic_aovs_miss(aov_normal_depth, aov_albedo_meshid, emit4);
    }
    else {
        const float pdfLight = ibl_pdf(emit, sc.cfg.avgIllum);
        misW = st->pdfb / (pdfLight + st->pdfb);
    }
    V3 contrib = misW * emit;
    contrib = contrib * st->throughput;
    st->contrib = st->contrib + contrib;
    st->terminated = true;
}

// aten::PathTracing::shade (src/libaten/renderer/pathtracing/pathtracing.cpp:91-236): the HOST
// order (normal flipped, then normal map), in three stages that hand over a small
// Surface record.  On the GPU each stage is its own kernel (the monolithic shade kernel was
// instruction-fetch bound); the CPU restatement simply calls them in a row (shade_hit below).
struct Surface {
    V3 p; float u;          // hit position, texture u
    V3 n; float v;          // orienting normal after flip + normal map, texture v
    V3 albedo; int32_t mtrlid;
};

// Stage 1: evaluate_hit, FillMaterial, albedo, HitTeminatedMaterial -> HitImplicitLight
// (pathtracing_impl.h:395-452), normal orientation + material::applyNormal.
// Returns false when the path ended here (implicit light).
// `svgf_w2c` != nullptr selects aten::SVGFRenderer::Shade (src/libaten/renderer/svgf/svgf.cpp:91-247) instead:
// the AOVs of the first hit (or of the hit behind a Specular first hit) take the geometric normal, the
// clip-space w of the hit as depth and the MATERIAL id in place of the mesh id; the albedo texture of that
// hit is taken out of the path (albedoMap = -1: "exporting separated albedo") and put back by the last
// a-trous pass; the albedo default is white instead of baseColor.
ATN_HD bool shade_surface(const SceneView& sc, const Ray& ray, int32_t objid, int32_t primid, float a, float b,
                          float t, int32_t mtrlid, int32_t bounce, PathState* st, Surface* sf,
                          AtnVec4* aov_normal_depth = nullptr, AtnVec4* aov_albedo_meshid = nullptr, int32_t meshid = -1,
                          const AtnMat4* svgf_w2c = nullptr)
{
    const AtnObjectParameter obj = sc.objects[objid];
    HitRecord rec;
    evaluate_hit(sc, obj, ray, primid, a, b, t, mtrlid, &rec);

    const bool backfacing = dot(rec.normal, -ray.dir) < 0.0f;
    V3 orienting = rec.normal;

    Mtrl mtrl = fill_material(sc, rec.mtrlid);
    AtnVec4 albedo4;
    if (svgf_w2c) {
        if ((bounce == 0 || (bounce == 1 && st->last_specular)) && aov_normal_depth && aov_albedo_meshid) {
            const AtnVec4 texcolor = sample_texture(sc, mtrl.albedoMap, rec.u, rec.v, v4(1.0f));
            fill_basic_aovs(aov_normal_depth, orienting, rec.p, *svgf_w2c, aov_albedo_meshid, texcolor, mtrlid);
            mtrl.albedoMap = -1;
        }
        albedo4 = sample_texture(sc, mtrl.albedoMap, rec.u, rec.v, v4(1.0f));
    }
    else {
        albedo4 = sample_texture(sc, mtrl.albedoMap, rec.u, rec.v, mtrl.baseColor);
    }

    if (!svgf_w2c && bounce == 0 && aov_normal_depth && aov_albedo_meshid) {
        // The reference's GPU shade kernel stores the AOVs of the first hit before anything else can end
        // the path, with the normal-mapped (not yet flipped) normal and an identity matrix
        // (pathtracing_impl.cu:92-111).
        AtnMat4 I;
        for (int i = 0; i < 16; i++) I.m[i] = (i % 5 == 0) ? 1.0f : 0.0f;
        const V3 aov_n = apply_normal_map(sc, mtrl.normalMap, rec.normal, rec.u, rec.v);
        fill_basic_aovs(aov_normal_depth, aov_n, rec.p, I, aov_albedo_meshid, albedo4, meshid);
    }

    if (mtrl.type == ATN_MTRL_EMISSIVE) {
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
            const V3 c = st->throughput * weight * light_color;
            st->contrib = st->contrib + c;
            st->terminated = true;
            return false;
        }
    }

    if (!mtrl.is_translucent() && backfacing) orienting = -orienting;
    orienting = apply_normal_map(sc, mtrl.normalMap, orienting, rec.u, rec.v);

    sf->p = rec.p; sf->u = rec.u; sf->n = orienting; sf->v = rec.v;
    sf->albedo = V3(albedo4); sf->mtrlid = rec.mtrlid;
    return true;
}

// Stage 2: FillShadowRay / SampleLight (pathtracing_impl.h:178-264) + ComputeRadianceNEE.
// In two steps, so that the GPU kernel can regroup its paths by the light that was picked before it samples it (k_nee):
// the pick (one random number) and the rest.  shade_nee() is the two in a row, which is what the CPU restatement runs.
ATN_HD int32_t shade_nee_pick(const SceneView& sc, uint32_t mtrl_attrib, Sampler* sampler)      // light index, or -1: no NEE
{
    const int32_t lightnum = sc.light_count;
    if (!(lightnum > 0) || (mtrl_attrib & (2u | 4u))) return -1;          // singular or translucent material
    int32_t target = (int32_t)(sampler->next() * lightnum);
    return (lightnum - 1 < target) ? (lightnum - 1) : target;
}
ATN_HD void shade_nee_light(const SceneView& sc, const V3& ray_dir, const Surface& sf, const Mtrl& mtrl, int32_t target,
                            PathState* st, ShadowRayOut* sh)
{
    const float select_prob = 1.0f / sc.light_count;
    const AtnLightParameter& light = sc.lights[target];
    LightSample ls;
    light_sample(&ls, light, sc, sf.p, sf.n, &st->sampler);

    sh->dir = normalize(ls.dir);
    sh->dist = length(ls.pos - sf.p);
    sh->org = ray_offset(sf.p, sf.n);
    sh->light_id = target;
    sh->lightcontrib = V3(0.0f);
    sh->sampled = true;
    V3 radiance;
    if (radiance_nee(&radiance, sc, ray_dir, sf.n, mtrl, sf.u, sf.v, select_prob, ls)) {
        sh->lightcontrib = st->throughput * radiance * sf.albedo;
        sh->active = true;
    }
}
ATN_HD void shade_nee(const SceneView& sc, const V3& ray_dir, const Surface& sf, PathState* st, ShadowRayOut* sh)
{
    sh->active = false;
    sh->sampled = false;
    const Mtrl mtrl = fill_material(sc, sf.mtrlid);
    const int32_t target = shade_nee_pick(sc, mtrl.attrib, &st->sampler);
    if (target >= 0) shade_nee_light(sc, ray_dir, sf, mtrl, target, st, sh);
}

// Stage 3: ComputeRussianProbability (:680-698), material::sampleMaterial, PrepareForNextBounce (:700-743).
ATN_HD void shade_bsdf(const SceneView& sc, const V3& ray_dir, const Surface& sf, int32_t bounce, int32_t rr_bounce,
                       PathState* st, Ray* next_ray)
{
    const Mtrl mtrl = fill_material(sc, sf.mtrlid);
    float russian_prob = 1.0f;
    if (bounce > rr_bounce) {
        if (squared_length(st->throughput) > 0) {
            russian_prob = max3(st->throughput);
            const float p = st->sampler.next();
            st->terminated = (p >= russian_prob);
        }
    }

    MtrlSample smp;
    smp.dir = V3(0.0f); smp.bsdf = V3(0.0f); smp.pdf = 0.0f;
    sample_material(&smp, sc, mtrl, sf.n, ray_dir, &st->sampler, sf.u, sf.v);

    const V3 next_dir = normalize(smp.dir);
    const float pdfb = smp.pdf;
    V3 along = sf.n;
    along = dot(along, next_dir) >= 0.0f ? along : -along;
    const float c = dot(along, next_dir);
    if (pdfb > 0 && c > 0) {
        st->throughput = st->throughput * (sf.albedo * smp.bsdf * c / pdfb);
        st->throughput = st->throughput / russian_prob;
    }
    else {
        st->terminated = true;
    }
    if (st->terminated) return;
    st->pdfb = pdfb;
    st->singular = mtrl.is_singular();
    // PrepareForNextBounce: last_hit_mtrl_idx = mtrl.id; SVGFRenderer::Shade asks whether that material is Specular
    st->last_specular = mtrl.id >= 0 && mtrl.id < sc.material_count && sc.materials[mtrl.id].type == ATN_MTRL_SPECULAR;
    *next_ray = make_ray(sf.p, next_dir, along);
}

ATN_HD void shade_hit(const SceneView& sc, const Ray& ray, int32_t objid, int32_t primid, float a, float b,
                      float t, int32_t mtrlid, int32_t bounce, int32_t rr_bounce,
                      PathState* st, ShadowRayOut* sh, Ray* next_ray,
                      AtnVec4* aov_normal_depth = nullptr, AtnVec4* aov_albedo_meshid = nullptr, const AtnMat4* svgf_w2c = nullptr)
{
    sh->active = false;
    if (st->terminated) return;
    Surface sf;
    if (!shade_surface(sc, ray, objid, primid, a, b, t, mtrlid, bounce, st, &sf, aov_normal_depth, aov_albedo_meshid, -1, svgf_w2c)) return;
    shade_nee(sc, ray.dir, sf, st, sh);
    shade_bsdf(sc, ray.dir, sf, bounce, rr_bounce, st, next_ray);
}

// ---------------------------------------------------------------------------------------------
// N1 (first renderer built on the path): ambient occlusion, AT_NAME::ao::ShandeByAO
// (src/libaten/renderer/ao/aorenderer_impl.h:33-117) and AORenderer::radiance
// (src/libaten/renderer/ao/aorenderer.cpp:20-58)
// ---------------------------------------------------------------------------------------------
// AORenderer::radiance re-seeds the sampler after GeneratePath (aorenderer.cpp:27-29)
ATN_HD Sampler ao_sampler_for_path(uint32_t frame, uint32_t rnd)
{
    Sampler s;
    s.scramble = rnd * 0x1fe3434fu * ((frame + 331u * rnd) / 256u);
    s.idx = frame % 256u;
    s.dim = 4 + 5 * 300;
    return s;
}
// First half of ShandeByAO: hit evaluation and the (normal-mapped, unflipped) shading normal.
ATN_HD void ao_surface(const SceneView& sc, const Ray& ray, int32_t objid, int32_t primid, float a, float b,
                       float t, int32_t mtrlid, V3* p, V3* n)
{
    HitRecord rec;
    evaluate_hit(sc, sc.objects[objid], ray, primid, a, b, t, mtrlid, &rec);
    const Mtrl m = fill_material(sc, rec.mtrlid);
    *p = rec.p;
    *n = apply_normal_map(sc, m.normalMap, rec.normal, rec.u, rec.v);
}
// One AO ray: Diffuse::sampleDirection + aten::ray(p, dir, normal) (aorenderer_impl.h:70-71)
ATN_HD Ray ao_ray(const V3& p, const V3& n, Sampler* sampler, V3* next_dir)
{
    const float r1 = sampler->next();
    const float r2 = sampler->next();
    *next_dir = lambert_dir(n, r1, r2);
    return make_ray(p, *next_dir, n);
}
// One lookup of the AO ray's loop (aorenderer_impl.h:72-111): the closest hit of `ray`, traced with t_max = radius.  A hit
// on a surface that is translucent by alpha is ignored: the ray restarts behind it (aten::ray(p, ray.dir, normal turned
// along the ray) -- direction normalised again, origin offset) and is traced again, up to 10 lookups; when the last one is
// ignored too, nothing is added.  Returns true when the ray has to be traced again from *next.
ATN_HD bool ao_accumulate(const SceneView& sc, float* ao, const V3& n, const V3& next_dir, const Ray& ray, float radius,
                          int32_t hit_objid, int32_t hit_primid, float hit_a, float hit_b, float hit_t, int32_t hit_mtrlid,
                          bool last_lookup, Ray* next)
{
    if (hit_objid < 0) { *ao = 1.0f; return false; }            // `ao_color = 1.0f` (assignment, as in the reference)
    HitRecord rec;
    evaluate_hit(sc, sc.objects[hit_objid], ray, hit_primid, hit_a, hit_b, hit_t, hit_mtrlid, &rec);
    if (hit_mtrlid >= 0) {                                      // material::isTranslucentByAlpha (material.cpp:193-210)
        const AtnMaterialParameter& m = sc.materials[hit_mtrlid];
        const float alpha = sample_texture(sc, m.albedoMap, rec.u, rec.v, v4(1.0f)).w * m.baseColor.w;
        if (alpha < 1.0f) {
            if (last_lookup) return false;
            V3 on = rec.normal;
            if (!(dot(rec.normal, ray.dir) > 0.0f)) on = -on;
            *next = make_ray(rec.p, ray.dir, on);
            return true;
        }
    }
    const float c = dot(rec.normal, next_dir);
    if (c > 0.0f) {
        const float pdfb = lambert_pdf(n, next_dir);
        *ao += hit_t / radius * c / pdfb;
    }
    return false;
}

// HitTestToTargetLight (pathtracing_impl.h:266-351) + scene::hitLight (src/libaten/scene/scene.h:120-133), one LOOKUP of
// its loop at a time.  The reference traces the shadow ray's closest hit (t_max = dist - 1e-9, with the ORIGINAL distance
// on every lookup); a hit on a surface that is translucent by alpha (material::isTranslucentByAlpha, material.cpp:193-210)
// or -- when the surface the shadow ray left has stencil type ALWAYS -- on a STENCIL surface is ignored: the ray restarts
// behind it (aten::ray(rec.p, original dir, normal turned along the ray): the direction is normalised AGAIN, the origin
// offset) and is traced again, up to shadow_max_lookups() times; when the last lookup is ignored too the light counts as
// not reached.  `hitobj` keeps the last object hit across lookups (the reference never resets it), so a lookup that
// finds nothing after an ignored hit does not reach an area light.
enum : int { kShadowHidden = 0, kShadowVisible = 1, kShadowRetrace = 2 };
enum : int { kStencilNone = 0, kStencilAlways = 1, kStencilStencil = 2 };       // aten::StencilType (material.h:224-228)

ATN_HD int32_t shadow_light_obj(const AtnLightParameter& light)
{
    return (light.type == 0 && light.arealight_objid >= 0) ? light.arealight_objid : -1;       // IsValidLightObjectId
}
// `surface_mtrlid`: the material of the surface the shadow ray starts from (HitShadowRay's surface_mtrl)
ATN_HD bool shadow_needs_stencil(const SceneView& sc, int32_t surface_mtrlid)
{
    return surface_mtrlid >= 0 && sc.materials[surface_mtrlid].stencil_type == kStencilAlways;
}
ATN_HD int shadow_max_lookups(const SceneView& sc, bool need_stencil) { return (sc.alpha_blending || need_stencil) ? 10 : 1; }

// `ray` is this lookup's ray, `orig_dir` the direction of the first one; `hit_*` its closest hit (hit_objid < 0: none);
// `*hitobj` starts as shadow_light_obj().  kShadowRetrace: trace again from *next_org along normalize(orig_dir).
ATN_HD int shadow_lookup(const SceneView& sc, const AtnLightParameter& light, const Ray& ray, const V3& orig_dir, float dist_to_light,
                         int32_t hit_objid, int32_t hit_primid, float hit_a, float hit_b, float hit_t, int32_t hit_mtrlid,
                         bool last_lookup, bool need_stencil, int32_t* hitobj, V3* next_org)
{
    const bool is_hit = hit_objid >= 0;
    if (is_hit) {
        *hitobj = hit_objid;
        const AtnMaterialParameter& m = sc.materials[hit_mtrlid < 0 ? 0 : hit_mtrlid];
        bool ignore = need_stencil && hit_mtrlid >= 0 && m.stencil_type == kStencilStencil;
        const bool textured = hit_mtrlid >= 0 && m.albedoMap >= 0;
        HitRecord rec;
        bool have_rec = false;
        float alpha = hit_mtrlid < 0 ? 1.0f : 1.0f * m.baseColor.w;
        if (textured) {
            evaluate_hit(sc, sc.objects[hit_objid], ray, hit_primid, hit_a, hit_b, hit_t, hit_mtrlid, &rec);
            have_rec = true;
            alpha = sample_texture(sc, m.albedoMap, rec.u, rec.v, v4(1.0f)).w * m.baseColor.w;
        }
        if (alpha < 1.0f) ignore = true;
        if (ignore) {
            if (last_lookup) return kShadowHidden;              // the loop runs out: is_hit_to_light stays false
            if (!have_rec) evaluate_hit(sc, sc.objects[hit_objid], ray, hit_primid, hit_a, hit_b, hit_t, hit_mtrlid, &rec);
            V3 n = rec.normal;
            if (!(dot(rec.normal, orig_dir) > 0.0f)) n = -n;    // "to go through the object, specify the opposite normal"
            *next_org = ray_offset(rec.p, n);
            return kShadowRetrace;
        }
    }
    if (*hitobj == shadow_light_obj(light)) return kShadowVisible;
    if (light.attrib & 2u) return is_hit ? kShadowHidden : kShadowVisible;
    if (light.attrib & 1u) return hit_t > dist_to_light ? kShadowVisible : kShadowHidden;
    return kShadowHidden;
}

// The single-lookup form (alpha blending off, no stencil): what every config of BASELINE.json runs.
ATN_HD bool shadow_visible(const SceneView& sc, const AtnLightParameter& light, const Ray& ray, float dist_to_light,
                           int32_t hit_objid, int32_t hit_primid, float hit_a, float hit_b, float hit_t, int32_t hit_mtrlid)
{
    int32_t hitobj = shadow_light_obj(light);
    V3 unused;
    return shadow_lookup(sc, light, ray, ray.dir, dist_to_light, hit_objid, hit_primid, hit_a, hit_b, hit_t, hit_mtrlid,
                         true, false, &hitobj, &unused) == kShadowVisible;
}

}  // namespace atn
