// atn_svgf.cuh -- the image-space half of the SVGF renderer (N1), one text for the sm_100a kernels
// (atn_kernels.cu) and for the CPU restatement under oracle/, which compiles this header with g++.
//
// Reference: aten::SVGFRenderer (src/libaten/renderer/svgf/svgf.cpp:248-639) and the per-pixel functions
// it shares with the device renderer idaten::SVGFPathTracing (src/libaten/renderer/svgf/svgf_impl.h).
// Frame k of the renderer is
//   path tracing, 1 sample (SVGFRenderer::Shade: separated albedo, AOVs)      svgf.cpp:20-247
//   PrepareForDenoise<k == 0>                                                   svgf_impl.h:112-141
//   k > 0: temporal reprojection + moment accumulation                          svgf_impl.h:204-366, svgf.cpp:248-307
//   variance estimation (7x7 / 5x5 bilateral while < 4 frames of history)       svgf_impl.h:438-544
//   5 a-trous wavelet iterations, albedo put back by the last one               svgf_impl.h:546-835, svgf.cpp:330-410
//   first a-trous result copied to the colour history                           svgf.cpp:412-424
// Every function takes pixel coordinates and plain AtnVec4 arrays (W*H, row major).
#pragma once

#include "atn_math.cuh"

#ifdef __CUDACC__
#define ATN_UNROLL _Pragma("unroll")
#else
#define ATN_UNROLL
#endif

namespace atn {

// color::luminance (src/libaten/misc/color.h:61-72)
ATN_HD float luminance(float r, float g, float b) { return 0.212639f * r + 0.71517f * g + 0.0721926f * b; }

// MatricesForRendering::Reset + GetW2C (src/libaten/renderer/pathtracing/pt_params.h:163-184):
// mat4::lookat(origin, center, up), mat4::perspective(znear, zfar, vfov, aspect) (mat4.h:457-515), V2C * W2V
// with mat4::operator*= (mat4.h:140-155).  Row-major m[r * 4 + c].
ATN_HD AtnMat4 camera_w2c(const AtnCameraParameter& cam)
{
    float w2v[16], v2c[16];
    for (int i = 0; i < 16; i++) { w2v[i] = (i % 5 == 0) ? 1.0f : 0.0f; v2c[i] = w2v[i]; }
    const V3 eye(cam.origin), at(cam.center), up(cam.up);
    const V3 zaxis = normalize(eye - at);
    const V3 xaxis = normalize(cross(up, zaxis));
    const V3 yaxis = cross(zaxis, xaxis);
    w2v[0] = xaxis.x; w2v[4] = yaxis.x; w2v[8] = zaxis.x;
    w2v[1] = xaxis.y; w2v[5] = yaxis.y; w2v[9] = zaxis.y;
    w2v[2] = xaxis.z; w2v[6] = yaxis.z; w2v[10] = zaxis.z;
    w2v[3] = -dot(xaxis, eye); w2v[7] = -dot(yaxis, eye); w2v[11] = -dot(zaxis, eye);
    w2v[15] = 1.0f;
    const float fH = 1 / tanf((kPi * (cam.vfov) / 180.0f) * 0.5f);
    const float fW = fH / cam.aspect;
    v2c[0] = fW; v2c[5] = fH;
    v2c[10] = cam.zfar / (cam.znear - cam.zfar);
    v2c[11] = cam.znear * cam.zfar / (cam.znear - cam.zfar);
    v2c[14] = -1.0f; v2c[15] = 0.0f;
    AtnMat4 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float t = 0.0f;
            for (int k = 0; k < 4; k++) t += v2c[i * 4 + k] * w2v[k * 4 + j];
            r.m[i * 4 + j] = t;
        }
    return r;
}

// Camera::ComputeScreenDistance (src/libaten/camera/camera.h:216-221); vfov goes in as it is (degrees)
ATN_HD float screen_distance(const AtnCameraParameter& cam, int32_t height) { return height / (2.0f * tanf(0.5f * cam.vfov)); }

// x^128 (the normal weight: pow(max(0, dot(n, n')), 128), svgf_impl.h:500,757).  The host build calls powf like
// the reference; the device build squares seven times -- 7 FMULs instead of the ~150-instruction powf that made
// the a-trous kernel issue bound -- which differs from the correctly rounded power by < 64 ulp (4e-6 relative),
// well inside the tolerance the GPU tests state for this float pipeline.
ATN_HD float svgf_pow128(float x)
{
#ifdef __CUDA_ARCH__
    float y = x * x; y *= y; y *= y; y *= y; y *= y; y *= y; y *= y;
    return y;
#else
    return powf(x, 128.0f);
#endif
}

ATN_HD int32_t svgf_clampi(int32_t v, int32_t lo, int32_t hi) { return v < lo ? lo : (v > hi ? hi : v); }
ATN_HD float svgf_clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

// One frame's AOV set (SVGFParams::aovs[k], svgf_types.h:55-66)
struct SvgfAov {
    AtnVec4* normal_depth;      // geometric normal of the first hit, clip-space w
    AtnVec4* albedo_meshid;     // albedo texel of the first hit, MATERIAL id (-1: background)
    AtnVec4* color_variance;    // filtered colour history, variance
    AtnVec4* moment_tw;         // sum lum^2, sum lum, frames accumulated, temporal weight
};

// PrepareForDenoise (svgf_impl.h:112-141).  `c` = PathContrib as vec4 (rgb, sample count).
ATN_HD void svgf_prepare(bool first_frame, int32_t idx, const AtnVec4& c, AtnVec4* temporary_color, const SvgfAov& aov)
{
    if (first_frame) {
        const float x = c.x / c.w, y = c.y / c.w, z = c.z / c.w;
        const float lum = luminance(x, y, z);
        AtnVec4 m = aov.moment_tw[idx];
        m.x += lum * lum; m.y += lum; m.z += 1;
        aov.moment_tw[idx] = m;
        AtnVec4 cv = aov.color_variance[idx];
        cv.x = x; cv.y = y; cv.z = z;
        aov.color_variance[idx] = cv;
    }
    temporary_color[idx] = c;
}

// SVGFRenderer::TemporalReprojection (svgf.cpp:248-307) = ExtractCenterPixel, UpdateAOVIfBackgroundPixel,
// svgf::TemporalReprojection (thresholds 0.98 / 0.05), AccumulateMoments.  Returns the colour the host
// renderer puts on the film at this point.  Every tap uses the motion vector of the CENTRE pixel
// (svgf_impl.h:317-318), as the reference does.
ATN_HD AtnVec4 svgf_temporal(int32_t ix, int32_t iy, int32_t width, int32_t height, float threshold_normal, float threshold_depth,
                             const AtnVec4& contrib, const SvgfAov& cur, const SvgfAov& prev, const AtnVec4* motion_depth)
{
    const int32_t idx = ix + iy * width;
    const AtnVec4 nd = cur.normal_depth[idx];
    const float center_depth = nd.w;
    const int32_t center_meshid = (int32_t)cur.albedo_meshid[idx].w;
    AtnVec4 curr; curr.x = contrib.x / contrib.w; curr.y = contrib.y / contrib.w; curr.z = contrib.z / contrib.w; curr.w = 1.0f / contrib.w;

    if (center_meshid < 0) {
        cur.color_variance[idx] = curr;
        AtnVec4 m = cur.moment_tw[idx];
        m.x = 1.0f; m.y = 1.0f; m.z = 1.0f;
        cur.moment_tw[idx] = m;
        return curr;
    }

    const V3 center_normal(nd.x, nd.y, nd.z);
    float sx = 0, sy = 0, sz = 0, sw = 0;
    float weight = 0.0f;
    const AtnVec4 md = motion_depth ? motion_depth[idx] : v4(0.0f);
    for (int32_t y = -1; y <= 1; y++) {
        for (int32_t x = -1; x <= 1; x++) {
            const int32_t xx = svgf_clampi(ix + x, 0, width - 1);
            const int32_t yy = svgf_clampi(iy + y, 0, height - 1);
            int32_t prev_x = (int32_t)(xx + md.x * width);
            int32_t prev_y = (int32_t)(yy + md.y * height);
            prev_x = svgf_clampi(prev_x, 0, width - 1);
            prev_y = svgf_clampi(prev_y, 0, height - 1);
            const int32_t pidx = prev_x + prev_y * width;
            const AtnVec4 pnd = prev.normal_depth[pidx];
            const float prev_depth = pnd.w;
            const int32_t prev_meshid = (int32_t)prev.albedo_meshid[pidx].w;
            const V3 prev_normal(pnd.x, pnd.y, pnd.z);
            const float Wz = svgf_clampf((threshold_depth - fabsf(1 - center_depth / prev_depth)) / threshold_depth, 0.0f, 1.0f);
            const float Wn = svgf_clampf((dot(center_normal, prev_normal) - threshold_normal) / (1.0f - threshold_normal), 0.0f, 1.0f);
            const float Wm = center_meshid == prev_meshid ? 1.0f : 0.0f;
            const AtnVec4 pc = prev.color_variance[pidx];
            const float W = Wz * Wn * Wm;
            sx += pc.x * W; sy += pc.y * W; sz += pc.z * W; sw += pc.w * W;
            weight += W;
        }
    }
    if (weight > 0.0f) {
        sx /= weight; sy /= weight; sz /= weight; sw /= weight;
        weight /= 9;
        curr.x = 0.2f * curr.x + 0.8f * sx; curr.y = 0.2f * curr.y + 0.8f * sy;
        curr.z = 0.2f * curr.z + 0.8f * sz; curr.w = 0.2f * curr.w + 0.8f * sw;
    }
    AtnVec4 m = cur.moment_tw[idx];
    m.w = weight;
    AtnVec4 cv = cur.color_variance[idx];
    cv.x = curr.x; cv.y = curr.y; cv.z = curr.z;
    cur.color_variance[idx] = cv;

    // AccumulateMoments (svgf_impl.h:204-250)
    const float lum = luminance(cv.x, cv.y, cv.z);
    float mx = lum * lum, my = lum;
    int32_t frame = 1;
    if (weight > 0.0f) {
        const AtnVec4 pm = prev.moment_tw[idx];
        frame = (int32_t)(pm.z + 1);
        mx += pm.x; my += pm.y;
    }
    m.x = mx; m.y = my; m.z = (float)frame;
    cur.moment_tw[idx] = m;
    return curr;
}

// svgf::EstimateVariance (svgf_impl.h:438-544).  The reference updates color_variance IN PLACE while its
// neighbours read it, so its result depends on the order in which pixels are processed (one host thread:
// row-major; several threads or the CUDA kernel: unspecified).  `cv_in` / `cv_out` make the order
// explicit: the same array reproduces the single-threaded host order when pixels are visited row-major;
// two arrays give the order-independent form the GPU path uses (every pixel reads the unfiltered
// colours).  Returns what the host renderer puts on the film (the variance).
ATN_HD AtnVec4 svgf_variance(int32_t ix, int32_t iy, int32_t width, int32_t height, float camera_distance,
                             const SvgfAov& aov, const AtnVec4* cv_in, AtnVec4* cv_out)
{
    const int32_t idx = ix + iy * width;
    const AtnVec4 nd = aov.normal_depth[idx];
    const AtnVec4 mt = aov.moment_tw[idx];
    const AtnVec4 center_color = cv_in[idx];
    const float center_depth = nd.w;
    const int32_t center_meshid = (int32_t)aov.albedo_meshid[idx].w;
    if (center_meshid < 0) {
        AtnVec4 m = mt; m.x = 0; m.y = 0; m.z = 1;
        aov.moment_tw[idx] = m;
        if (cv_out != cv_in) cv_out[idx] = center_color;
        return v4(0.0f);
    }
    const float pixel_distance_ratio = (center_depth / camera_distance) * height;
    const int32_t frame = (int32_t)mt.z;
    float cmx = mt.x / mt.z, cmy = mt.y / mt.z, cmz = mt.z / mt.z;
    float variance = 0.0f;
    AtnVec4 color = center_color;
    if (frame < 4) {
        const V3 center_normal(nd.x, nd.y, nd.z);
        float msx = cmx, msy = cmy, msz = cmz;
        float weight = 1.0f;
        const int32_t radius = frame > 1 ? 2 : 3;
        for (int32_t v = -radius; v <= radius; v++) {
            for (int32_t u = -radius; u <= radius; u++) {
                if (u != 0 || v != 0) {
                    const int32_t xx = svgf_clampi(ix + u, 0, width - 1);
                    const int32_t yy = svgf_clampi(iy + v, 0, height - 1);
                    const int32_t sidx = xx + yy * width;
                    const AtnVec4 snd = aov.normal_depth[sidx];
                    const AtnVec4 smt = aov.moment_tw[sidx];
                    const V3 sample_nml(snd.x, snd.y, snd.z);
                    const float sample_depth = snd.w;
                    const int32_t sample_meshid = (int32_t)aov.albedo_meshid[sidx].w;
                    const AtnVec4 sample_color = cv_in[sidx];
                    const float mx = smt.x / smt.z, my = smt.y / smt.z, mz = smt.z / smt.z;
                    const float uv_length = sqrtf((float)(u * u + v * v));
                    const float Wz = fabsf(sample_depth - center_depth) / (pixel_distance_ratio * uv_length + 1e-2f);
                    const float Wn = svgf_pow128(fmaxf(0.0f, dot(sample_nml, center_normal)));
                    const float Wm = center_meshid == sample_meshid ? 1.0f : 0.0f;
                    const float W = expf(-Wz) * Wn * Wm;
                    msx += mx * W; msy += my * W; msz += mz * W;
                    color.x += sample_color.x * W; color.y += sample_color.y * W; color.z += sample_color.z * W; color.w += sample_color.w * W;
                    weight += W;
                }
            }
        }
        msx /= weight; msy /= weight;
        color.x /= weight; color.y /= weight; color.z /= weight; color.w /= weight;
        variance = fmaxf(0.0f, msx - msy * msy);
    }
    else {
        variance = fmaxf(0.0f, cmx - cmy * cmy);
    }
    color.w = variance;
    cv_out[idx] = color;
    AtnVec4 r; r.x = variance; r.y = variance; r.z = variance; r.w = 1.0f;
    return r;
}

// SVGFRenderer::AtrousFilter (svgf.cpp:330-410): ExtractCenterPixel<false>, CheckIfBackgroundPixelForAtrous,
// Exec3x3GaussFilter of the variance, ExecAtrousWaveletFilter, PostProcessForAtrousFilter.
// `src` = aov.color_variance for the first iteration, the ping-pong buffer after that.  Returns true and
// the film colour when this iteration produces one (the last iteration; albedo put back).
// `tap_nd(i)` / `tap_color(i)` read the normal-depth plane and the colour source at pixel i for the 24 taps: plain
// array reads on the host (SvgfPlane); the kernel passes readers that go through the texture pipe for one of them
// (the a-trous kernel is bound by the LSU pipe of L1, see k_svgf_atrous).
struct SvgfPlane {
    const AtnVec4* p;
    ATN_HD AtnVec4 operator()(int32_t i) const { return p[i]; }
};
template <class TapNd, class TapColor>
ATN_HD bool svgf_atrous(int32_t iter, int32_t iter_count, int32_t ix, int32_t iy, int32_t width, int32_t height,
                        float camera_distance, const SvgfAov& aov, const AtnVec4* src, AtnVec4* next,
                        AtnVec4* temporary_color, AtnVec4* film_color, const TapNd& tap_nd, const TapColor& tap_color)
{
    const bool first = iter == 0, last = iter == iter_count - 1;
    const int32_t idx = ix + iy * width;
    const AtnVec4 nd = aov.normal_depth[idx];
    const AtnVec4 am = aov.albedo_meshid[idx];
    const float center_depth = nd.w;
    const int32_t center_meshid = (int32_t)am.w;
    const AtnVec4 s = src[idx];
    AtnVec4 center_color; center_color.x = s.x; center_color.y = s.y; center_color.z = s.z; center_color.w = 1.0f;
    const V3 center_normal(nd.x, nd.y, nd.z);

    if (am.w < 0) {      // the reference compares the float here (svgf_impl.h:627-629)
        AtnVec4 o; o.x = center_color.x; o.y = center_color.y; o.z = center_color.z; o.w = 0.0f;
        next[idx] = o;
        if (last) {
            film_color->x = am.x * center_color.x; film_color->y = am.y * center_color.y;
            film_color->z = am.z * center_color.z; film_color->w = am.w * center_color.w;
            return true;
        }
        // Not the last iteration: CheckIfBackgroundPixelForAtrous returns an engaged optional that holds an
        // EMPTY optional, which SVGFRenderer::AtrousFilter (svgf.cpp:365-371) does not treat as "done" -- the
        // background pixel goes through the filter below like any other (all its taps weigh 0 because its
        // normal is 0, so it keeps its colour and gets variance 1).  Reproduced, because that variance
        // feeds the 3x3 Gauss filter of the pixels next to the background.
    }

    // 3x3 Gauss filter of the variance channel (svgf_impl.h:546-603)
    float gauss = 0;
    {
        static constexpr float kernel[9] = { 1.0f / 16.0f, 1.0f / 8.0f, 1.0f / 16.0f, 1.0f / 8.0f, 1.0f / 4.0f, 1.0f / 8.0f,
                                  1.0f / 16.0f, 1.0f / 8.0f, 1.0f / 16.0f };
        int pos = 0;
        ATN_UNROLL
        for (int32_t oy = -1; oy <= 1; oy++)
            ATN_UNROLL
            for (int32_t ox = -1; ox <= 1; ox++) {
                const int32_t xx = svgf_clampi(ix + ox, 0, width - 1);
                const int32_t yy = svgf_clampi(iy + oy, 0, height - 1);
                gauss += kernel[pos++] * src[xx + yy * width].w;
            }
    }

    // ExecAtrousWaveletFilter (svgf_impl.h:665-779)
    const float sigmaZ = 1.0f, sigmaL = 4.0f;
    static constexpr float h[24] = { 2.0f / 3.0f, 2.0f / 3.0f, 2.0f / 3.0f, 2.0f / 3.0f, 1.0f / 6.0f, 1.0f / 6.0f, 1.0f / 6.0f, 1.0f / 6.0f,
                          4.0f / 9.0f, 4.0f / 9.0f, 4.0f / 9.0f, 4.0f / 9.0f, 1.0f / 9.0f, 1.0f / 9.0f, 1.0f / 9.0f, 1.0f / 9.0f,
                          1.0f / 9.0f, 1.0f / 9.0f, 1.0f / 9.0f, 1.0f / 9.0f, 1.0f / 36.0f, 1.0f / 36.0f, 1.0f / 36.0f, 1.0f / 36.0f };
    static constexpr int8_t offx[24] = { 1, 0, -1, 0, 2, 0, -2, 0, 1, -1, -1, 1, 1, -1, -1, 1, 2, -2, -2, 2, 2, -2, -2, 2 };
    static constexpr int8_t offy[24] = { 0, 1, 0, -1, 0, 2, 0, -2, 1, 1, -1, -1, 2, 2, -2, -2, 1, 1, -1, -1, 2, 2, -2, -2 };
    const int32_t step_scale = 1 << iter;
    const float sqrt_gauss = sqrtf(gauss);
    const float center_lum = luminance(center_color.x, center_color.y, center_color.z);
    float cx = center_color.x, cy = center_color.y, cz = center_color.z, cw = center_color.w;
    float sumV = center_color.w;
    float weight = 1.0f;
    const float pixel_distance_ratio = (center_depth / camera_distance) * height;
    ATN_UNROLL
    for (int i = 0; i < 24; i++) {
        const int32_t sox = offx[i] * step_scale, soy = offy[i] * step_scale;
        const int32_t xx = svgf_clampi(ix + sox, 0, width - 1);
        const int32_t yy = svgf_clampi(iy + soy, 0, height - 1);
        const float u_length = sqrtf((float)(sox * sox + soy * soy));
        const int32_t qidx = xx + yy * width;
        const AtnVec4 qnd = tap_nd(qidx);
        const V3 normal(qnd.x, qnd.y, qnd.z);
        const float depth = qnd.w;
        const int32_t meshid = (int32_t)aov.albedo_meshid[qidx].w;
        const AtnVec4 color = tap_color(qidx);      // = aov.color_variance in the first iteration, the ping-pong source after
        const float variance = color.w;
        const float lum = luminance(color.x, color.y, color.z);
        const float Wn = svgf_pow128(fmaxf(0.0f, dot(center_normal, normal)));     // sigmaN = 128
#ifdef __CUDA_ARCH__
        const float Wz = 3.0f * fabsf(center_depth - depth) / (sigmaZ * (pixel_distance_ratio * u_length) + 0.000001f);
        const float Wl = fminf(expf(-fabsf(center_lum - lum) / (sigmaL * sqrt_gauss + 0.000001f)), 1.0f);
#else
        // the reference's host build resolves the unqualified fabs() in these two lines to the C double overload
        // (svgf_impl.h:755,759), so the quotient is formed in double and rounded once; the device build rounds in float
        const float Wz = (float)(3.0f * fabs((double)(center_depth - depth)) / (double)(sigmaZ * (pixel_distance_ratio * u_length) + 0.000001f));
        const float Wl = fminf(expf((float)(-fabs((double)(center_lum - lum)) / (double)(sigmaL * sqrt_gauss + 0.000001f))), 1.0f);
#endif
        const float Wm = meshid == center_meshid ? 1.0f : 0.0f;
        const float W = expf(-Wl * Wl - Wz) * Wn * Wm * h[i];
        cx += W * color.x; cy += W * color.y; cz += W * color.z; cw += W * color.w;
        sumV += W * W * variance;
        weight += W;
    }
    cx /= weight; cy /= weight; cz /= weight;
    sumV /= (weight * weight);
    AtnVec4 filtered; filtered.x = cx; filtered.y = cy; filtered.z = cz; filtered.w = sumV;

    // PostProcessForAtrousFilter (svgf_impl.h:795-823)
    next[idx] = filtered;
    if (first) {
        AtnVec4 t = temporary_color[idx];
        t.x = filtered.x; t.y = filtered.y; t.z = filtered.z;
        temporary_color[idx] = t;
    }
    if (last) {
        film_color->x = am.x * filtered.x; film_color->y = am.y * filtered.y;
        film_color->z = am.z * filtered.z; film_color->w = am.w * filtered.w;
        return true;
    }
    return false;
}

ATN_HD bool svgf_atrous(int32_t iter, int32_t iter_count, int32_t ix, int32_t iy, int32_t width, int32_t height,
                        float camera_distance, const SvgfAov& aov, const AtnVec4* src, AtnVec4* next,
                        AtnVec4* temporary_color, AtnVec4* film_color)
{
    return svgf_atrous(iter, iter_count, ix, iy, width, height, camera_distance, aov, src, next, temporary_color, film_color,
                       SvgfPlane{ aov.normal_depth }, SvgfPlane{ iter == 0 ? aov.color_variance : src });
}

// CopyFromTeporaryColorBufferToAov (svgf.cpp:412-424): xyz only
ATN_HD void svgf_copy(int32_t idx, const AtnVec4* temporary_color, const SvgfAov& aov)
{
    const AtnVec4 t = temporary_color[idx];
    AtnVec4 d = aov.color_variance[idx];
    d.x = t.x; d.y = t.y; d.z = t.z;
    aov.color_variance[idx] = d;
}

}  // namespace atn
