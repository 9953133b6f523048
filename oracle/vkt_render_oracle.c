/* TEST INFRASTRUCTURE ONLY -- part of oracle/libvkt_oracle.so.  Never linked into or imported by the product.
 *
 * Plain-C restatement of the reference's headless ray-marcher: RayMarchingKernel (absorption + emission) of
 * src/vkt/Render_kernel.hpp:95-162 as the reference's host path configures it (src/vkt/Render.cpp:434-489), driven by
 * visionaray's pinhole camera with the pixel-centre sampler.  The templates it restates are vendored with the
 * reference under src/3rdparty/visionaray/include/visionaray (cited as vsn/...).  PINNED: oracle/ref_render_shim.cpp
 * compiles the reference's own kernel and visionaray's own camera, and tests/test_render.py compares this file with
 * it float for float.  Binary32 arithmetic, one rounding per operation (-ffp-contract=off), same libm calls.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "vkt_oracle.h"

typedef struct { float x, y, z; } v3;

static v3 v3_(float x, float y, float z) { v3 r = { x, y, z }; return r; }
static v3 sub(v3 a, v3 b) { return v3_(a.x - b.x, a.y - b.y, a.z - b.z); }
static v3 add(v3 a, v3 b) { return v3_(a.x + b.x, a.y + b.y, a.z + b.z); }
static v3 scale(v3 a, float s) { return v3_(a.x * s, a.y * s, a.z * s); }
static v3 divv(v3 a, v3 b) { return v3_(a.x / b.x, a.y / b.y, a.z / b.z); }
/* vsn/math/detail/vector3.inl:292-307, 332-335; rsqrt = 1 / sqrt on the host (vsn/math/detail/math.h:537-542) */
static v3 cross(v3 u, v3 v) { return v3_(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x); }
static float dot(v3 u, v3 v) { return u.x * v.x + u.y * v.y + u.z * v.z; }
static v3 normalize(v3 v) { return scale(v, 1.0f / sqrtf(dot(v, v))); }
/* vsn/math/detail/math.h:88-98, 500-503 */
static float vmin(float x, float y) { return x < y ? x : y; }
static float vmax(float x, float y) { return x < y ? y : x; }
static float vclamp(float x, float a, float b) { return vmax(a, vmin(x, b)); }

/* (float)tex3D(volume, coord): Nearest filter, Clamp addressing, normalised coordinates
 * (vsn/texture/detail/texture_common.h:115-119, filter/nearest.h:87-104) */
static float fetch(const vko_vol* v, v3 c)
{
    const float N[3] = { (float)v->dims[0], (float)v->dims[1], (float)v->dims[2] };
    int i[3];
    const float cc[3] = { c.x, c.y, c.z };
    for (int d = 0; d < 3; ++d)
        i[d] = (int)(vclamp(cc[d], 0.0f, 1.0f - 1.0f / N[d]) * N[d]);
    const size_t idx = ((size_t)i[2] * (size_t)v->dims[1] + (size_t)i[1]) * (size_t)v->dims[0] + (size_t)i[0];
    switch (v->format) {
    case VKO_UINT8:  return (float)((const uint8_t*)v->data)[idx];
    case VKO_UINT16: return (float)((const uint16_t*)v->data)[idx];
    case VKO_INT16:  return (float)((const int16_t*)v->data)[idx];
    case VKO_UINT32: return (float)((const uint32_t*)v->data)[idx];
    }
    return 0.f;
}

/* normalize(Volume, voxel): (voxel - min) / (max - min) of the texel type (Render_kernel.hpp:31-39) */
static float normalize_voxel(int format, float voxel)
{
    switch (format) {
    case VKO_UINT8:  voxel -= 0.f; voxel /= (float)(255 - 0); break;
    case VKO_UINT16: voxel -= 0.f; voxel /= (float)(65535 - 0); break;
    case VKO_INT16:  voxel -= (float)-32768; voxel /= (float)(32767 - -32768); break;
    case VKO_UINT32: voxel -= 0.f; voxel /= (float)(4294967295u - 0u); break;
    }
    return voxel;
}

/* linear_to_srgb, vsn/detail/color_conversion.h:236-250 */
static float to_srgb(float c)
{
    return c <= 0.0031308f ? 12.92f * c : 1.055f * powf(c, (float)(1.0 / 2.4)) - 0.055f;
}

int vko_render_supported_format(int format)
{
    return format == VKO_UINT8 || format == VKO_UINT16 || format == VKO_INT16 || format == VKO_UINT32;
}

int vko_render_raymarch(const vko_vol* vol, const float* lut, int lut_n, const float eye_[3], const float center_[3],
                        const float up_[3], float fovy_degrees, int w, int h, float dt, int srgb, const float dist[3],
                        float* rgba32f, uint8_t* rgba8)
{
    if (!vko_render_supported_format(vol->format) || w <= 0 || h <= 0) return -1;
    /* pinhole_camera::perspective / begin_frame (vsn/detail/pinhole_camera.inl:31-38, 83-94) */
    const float fovy = fovy_degrees * (float)(3.14159265358979323846 / 180.0);
    const float aspect = (float)w / (float)h;
    const v3 eye = v3_(eye_[0], eye_[1], eye_[2]), center = v3_(center_[0], center_[1], center_[2]);
    const v3 up = v3_(up_[0], up_[1], up_[2]);
    const v3 f = normalize(sub(eye, center));
    const v3 s = normalize(cross(up, f));
    const v3 u = cross(f, s);
    const v3 U = scale(scale(s, tanf(fovy / 2.0f)), aspect);
    const v3 V = scale(u, tanf(fovy / 2.0f));
    const v3 W = v3_(-f.x, -f.y, -f.z);
    /* bbox = [0, dims * dist] (Render.cpp:838-841); size() = max - min */
    const v3 bmin = v3_(0.f, 0.f, 0.f);
    const v3 bmax = v3_((float)vol->dims[0] * dist[0], (float)vol->dims[1] * dist[1], (float)vol->dims[2] * dist[2]);
    const v3 box = sub(bmax, bmin);
    const float width = (float)w, height = (float)h;

    for (int py = 0; py < h; ++py)
        for (int px = 0; px < w; ++px) {
            /* pinhole_camera::primary_ray, vsn/detail/pinhole_camera.inl:103-115 (pixel-centre sampler: offset 0) */
            const float cu = 2.0f * ((float)px + 0.5f) / width - 1.0f;
            const float cv = 2.0f * ((float)py + 0.5f) / height - 1.0f;
            const v3 ori = eye;
            const v3 dir = normalize(add(add(scale(U, cu), scale(V, cv)), W));
            /* intersect(ray, aabb), vsn/math/intersect.h:43-72 */
            const v3 inv = v3_(1.0f / dir.x, 1.0f / dir.y, 1.0f / dir.z);
            const v3 t1 = v3_((bmin.x - ori.x) * inv.x, (bmin.y - ori.y) * inv.y, (bmin.z - ori.z) * inv.z);
            const v3 t2 = v3_((bmax.x - ori.x) * inv.x, (bmax.y - ori.y) * inv.y, (bmax.z - ori.z) * inv.z);
            const float tnear = vmax(vmin(t1.x, t2.x), vmax(vmin(t1.y, t2.y), vmin(t1.z, t2.z)));
            const float tfar = vmin(vmax(t1.x, t2.x), vmin(vmax(t1.y, t2.y), vmax(t1.z, t2.z)));
            /* RayMarchingKernel::operator(), Render_kernel.hpp:100-158 */
            float t = tnear;
            const v3 pos = add(ori, scale(dir, t));
            v3 tc = divv(pos, box);
            const v3 inc = divv(scale(dir, dt), box);
            float dst[4] = { 0.f, 0.f, 0.f, 0.f };
            while (t < tfar) {
                float voxel = fetch(vol, tc);
                voxel = normalize_voxel(vol->format, voxel);
                float color[4];
                if (lut && lut_n > 0) {
                    /* tex1D(transfunc, voxel): Nearest, Clamp (filter/nearest.h:29-45) */
                    const float n = (float)lut_n;
                    const int i = (int)(vclamp(voxel, 0.0f, 1.0f - 1.0f / n) * n);
                    memcpy(color, lut + 4 * (size_t)i, sizeof color);
                } else {
                    color[0] = color[1] = color[2] = color[3] = voxel;
                }
                color[3] = 1.f - powf(1.f - color[3], dt);         /* opacity correction */
                color[0] *= color[3]; color[1] *= color[3]; color[2] *= color[3];
                const float k = 1.0f - dst[3];                      /* front-to-back compositing; t < tfar holds here */
                for (int c = 0; c < 4; ++c) dst[c] += color[c] * k;
                /* the early-ray termination of :143 tests result.color, which is still 0: it never fires */
                tc = add(tc, inc);
                t += dt;
            }
            /* AccumulationKernel::accum with frameNum = 1 and an empty buffer (:60-80): (1 - 1) * 0 + 1 * src */
            float out[4];
            for (int c = 0; c < 4; ++c) out[c] = (1.f - 1.f) * 0.f + 1.f * dst[c];
            if (srgb) { out[0] = to_srgb(out[0]); out[1] = to_srgb(out[1]); out[2] = to_srgb(out[2]); }
            float* o = rgba32f + ((size_t)py * (size_t)w + (size_t)px) * 4;
            memcpy(o, out, sizeof out);
            if (rgba8)
                for (int c = 0; c < 4; ++c) {
                    /* float -> unorm<8>: saturate, multiply by 255 in binary64, truncate (vsn/math/detail/unorm.inl:20-24) */
                    const float sat = vmax(0.0f, vmin(out[c], 1.0f));
                    rgba8[((size_t)py * (size_t)w + (size_t)px) * 4 + c] = (uint8_t)(uint32_t)((double)sat * 255.0);
                }
        }
    return 0;
}

/* the viewer's default camera (Render.cpp:856-865; pinhole_camera::view_all, vsn/detail/pinhole_camera.inl:70-80) */
void vko_render_default_camera(const int32_t dims[3], const float dist[3], float eye[3], float center[3], float up[3])
{
    const float fovy = 45.f * (float)(3.14159265358979323846 / 180.0);
    const v3 size = v3_((float)dims[0] * dist[0], (float)dims[1] * dist[1], (float)dims[2] * dist[2]);
    const float diagonal = sqrtf(dot(size, size));
    const float r = diagonal * 0.5f;
    /* box.center() = (min + max) * 0.5 */
    center[0] = size.x * 0.5f; center[1] = size.y * 0.5f; center[2] = size.z * 0.5f;
    eye[0] = center[0]; eye[1] = center[1]; eye[2] = center[2] + (r + r / atanf(fovy));
    up[0] = 0.f; up[1] = 1.f; up[2] = 0.f;
}
