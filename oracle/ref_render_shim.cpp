// TEST INFRASTRUCTURE ONLY -- part of oracle/_ref/libvkt_ref_render.so.
//
// Runs the UNMODIFIED ray-marching kernel of the reference (src/vkt/Render_kernel.hpp:95-162, RayMarchingKernel) on
// the CPU exactly as the reference's host path sets it up (src/vkt/Render.cpp:434-489: volume texture with Nearest
// filter and Clamp addressing over normalised coordinates, RGBA look-up table as a Nearest 1-D texture, bbox =
// [0, dims * dist]), one frame (frameNum = 1, empty accumulation buffer).  What the reference does around the kernel
// needs a window (GLUT viewer, jittered progressive accumulation): the headless form drives the kernel with
// visionaray's own pinhole_camera and the pixel-centre sampler (pixel_sampler::uniform_type), i.e.
// cam.primary_ray(x, y, width, height) -- no random numbers, so an image is reproducible.
#include <cstdint>
#include <cstring>
#include <vector>

#include <visionaray/detail/macros.h>
#include <visionaray/detail/color_conversion.h>
#include <visionaray/math/math.h>
#include <visionaray/math/unorm.h>
#include <visionaray/pinhole_camera.h>
#include <visionaray/random_generator.h>
#include <visionaray/texture/texture.h>

#include <vkt/common.hpp>

#include "Render_kernel.hpp"

using namespace visionaray;

namespace {

struct RefVol { void* data; int32_t dims[3]; int32_t format; float lo, hi; };

template <typename Texel>
int render(RefVol const* vol, float const* lut, int lutN, pinhole_camera& cam, int w, int h, float dt, int srgb,
           float const* dist, float* rgba32f, uint8_t* rgba8)
{
    using Vol = texture_ref<Texel, 3>;
    using TF = texture_ref<vec4, 1>;
    Vol tex(vol->dims[0], vol->dims[1], vol->dims[2]);
    tex.reset((Texel*)vol->data);
    tex.set_filter_mode(Nearest);
    tex.set_address_mode(Clamp);
    TF tf(0U);
    if (lut && lutN > 0) {
        tf = TF((unsigned)lutN);
        tf.set_filter_mode(Nearest);
        tf.set_address_mode(Clamp);
        tf.reset((vec4*)lut);
    }
    std::vector<vec4f> accum((size_t)w * h, vec4f(0.f));
    RayMarchingKernel<Vol, TF> kernel;
    kernel.bbox = aabb(vec3(0.f), vec3(vol->dims[0] * dist[0], vol->dims[1] * dist[1], vol->dims[2] * dist[2]));
    kernel.volume = tex;
    kernel.transfunc = tf;
    kernel.dt = dt;
    kernel.width = w;
    kernel.height = h;
    kernel.frameNum = 1;
    kernel.accumBuffer = accum.data();
    kernel.sRGB = srgb != 0;
    random_generator<float> gen(0);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            auto ray = cam.primary_ray(basic_ray<float>{}, (float)x, (float)y, (float)w, (float)h);
            auto res = kernel(ray, gen, x, y);
            float* o = rgba32f + ((size_t)y * w + x) * 4;
            o[0] = res.color.x; o[1] = res.color.y; o[2] = res.color.z; o[3] = res.color.w;
            if (rgba8)
                for (int c = 0; c < 4; ++c)       // visionaray's own float -> unorm<8> rule (math/detail/unorm.inl:20-24)
                    rgba8[((size_t)y * w + x) * 4 + c] = unorm<8>(o[c]).value;
        }
    return 0;
}

} // namespace

extern "C" int ref_render_raymarch(RefVol const* vol, float const* lut, int lutN, float const* eye, float const* center,
                                   float const* up, float fovyDegrees, int w, int h, float dt, int srgb, float const* dist,
                                   float* rgba32f, uint8_t* rgba8)
{
    pinhole_camera cam;
    cam.look_at(vec3(eye[0], eye[1], eye[2]), vec3(center[0], center[1], center[2]), vec3(up[0], up[1], up[2]));
    cam.perspective(fovyDegrees * constants::degrees_to_radians<float>(), w / static_cast<float>(h), .001f, 1000.f);
    cam.set_viewport(0, 0, w, h);
    cam.begin_frame();
    switch ((vkt::DataFormat)vol->format) {
    case vkt::DataFormat::UInt8:  return render<uint8_t>(vol, lut, lutN, cam, w, h, dt, srgb, dist, rgba32f, rgba8);
    case vkt::DataFormat::UInt16: return render<uint16_t>(vol, lut, lutN, cam, w, h, dt, srgb, dist, rgba32f, rgba8);
    case vkt::DataFormat::Int16:  return render<int16_t>(vol, lut, lutN, cam, w, h, dt, srgb, dist, rgba32f, rgba8);
    case vkt::DataFormat::UInt32: return render<uint32_t>(vol, lut, lutN, cam, w, h, dt, srgb, dist, rgba32f, rgba8);
    default: return -1;   // Float32: normalize() divides by numeric_limits<float>::max() - min(): not meaningful
    }
}

// the default camera of the viewer (Render.cpp:856-865): 45 degrees, view_all(bbox); returns eye / center / up
extern "C" void ref_render_default_camera(int32_t const* dims, float const* dist, int w, int h, float* eye, float* center, float* up)
{
    pinhole_camera cam;
    cam.perspective(45.f * constants::degrees_to_radians<float>(), w / static_cast<float>(h), .001f, 1000.f);
    cam.view_all(aabb(vec3(0.f), vec3(dims[0] * dist[0], dims[1] * dist[1], dims[2] * dist[2])));
    for (int i = 0; i < 3; ++i) { eye[i] = cam.eye()[i]; center[i] = cam.center()[i]; up[i] = cam.up()[i]; }
}
