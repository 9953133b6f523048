// Headless ray-marcher (last row): the arithmetic of the reference's RayMarchingKernel
// (src/vkt/Render_kernel.hpp:95-162: absorption + emission, front-to-back compositing) as a CUDA kernel that writes
// RGBA into a device buffer.  No window, no GL: what the reference puts around the kernel (GLUT viewer, progressive
// jittered accumulation, src/vkt/Render.cpp) is replaced by ONE frame of pixel-centre rays of a pinhole camera.
//
// Set-up as the reference's host path (Render.cpp:434-489): the volume is a Nearest-filtered, Clamp-addressed texture
// over normalised coordinates, the RGBA look-up table a Nearest 1-D texture, bbox = [0, dims * dist].  Every float
// operation is the reference's, in its order, rounded on its own (the library is built with --fmad=false; divisions
// and square roots use the IEEE intrinsics), so with dt == 1 (pow(x, 1) == x) and sRGB off the float image equals the
// CPU's bit for bit; powf (opacity correction with dt != 1, sRGB) differs from libm by an ulp or two, which is where
// the +-1 LSB tolerance on RGBA8 comes from.
//
// Mapping: a warp owns an 8 x 4 pixel tile (neighbouring rays march through neighbouring voxels, so a warp's 32
// fetches of one step fall into few 32-byte sectors), a CTA 4 x 2 such tiles; one thread per pixel marches its ray.
// Traffic: every step reads one voxel through the read-only path; a ray touches ~ (tfar - tnear) / dt voxels.
#include "common.cuh"
#include "launch.hpp"

#include <cmath>

namespace vktb {

namespace {

struct RenderParams
{
    const uint8_t* vol;
    int dimX, dimY, dimZ;
    int format;
    float nX, nY, nZ;            // (float)dims
    float cX, cY, cZ;            // 1 - 1 / (float)dims: upper clamp of the normalised coordinate
    float eye[3], U[3], V[3], W[3];
    float bmax[3];               // bbox max (min = 0); size() = max - min = max
    const float4* lut;
    int lutN;
    float lutNf, lutC;           // (float)lutN, 1 - 1 / (float)lutN
    float dt;
    int srgb;
    int width, height;
    float widthF, heightF;
    uint8_t* rgba8;
    float4* rgba32f;
};

__device__ __forceinline__ float refMin(float x, float y) { return x < y ? x : y; }      // visionaray min / max
__device__ __forceinline__ float refMax(float x, float y) { return x < y ? y : x; }
__device__ __forceinline__ float refClamp(float x, float a, float b) { return refMax(a, refMin(x, b)); }

template <int F> __device__ __forceinline__ float fetchVoxel(const RenderParams& p, size_t idx)
{
    if (F == F_U8) return (float)__ldg(p.vol + idx);
    if (F == F_U16) return (float)__ldg(reinterpret_cast<const unsigned short*>(p.vol) + idx);
    if (F == F_INT16) return (float)__ldg(reinterpret_cast<const short*>(p.vol) + idx);
    return (float)__ldg(reinterpret_cast<const unsigned*>(p.vol) + idx);
}

template <int F> __device__ __forceinline__ float normalizeVoxel(float v)
{
    // (voxel - min) / (max - min) of the texel type, Render_kernel.hpp:31-39
    if (F == F_U8) return __fdiv_rn(__fsub_rn(v, 0.f), 255.f);
    if (F == F_U16) return __fdiv_rn(__fsub_rn(v, 0.f), 65535.f);
    if (F == F_INT16) return __fdiv_rn(__fsub_rn(v, -32768.f), 65535.f);
    return __fdiv_rn(__fsub_rn(v, 0.f), 4294967296.f);      // (float)4294967295u
}

__device__ __forceinline__ float toSrgb(float c)
{
    return c <= 0.0031308f ? __fmul_rn(12.92f, c) : __fsub_rn(__fmul_rn(1.055f, powf(c, (float)(1.0 / 2.4))), 0.055f);
}

template <int F, bool LUT>
__global__ void __launch_bounds__(256)
rayMarchKernel(RenderParams p)
{
    // warp = 8 x 4 pixels, CTA = 4 x 2 warps = 32 x 8 pixels
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int px = blockIdx.x * 32 + (warp & 3) * 8 + (lane & 7);
    const int py = blockIdx.y * 8 + (warp >> 2) * 4 + (lane >> 3);
    if (px >= p.width || py >= p.height) return;

    // pinhole_camera::primary_ray (pixel centre)
    const float cu = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, __fadd_rn((float)px, 0.5f)), p.widthF), 1.0f);
    const float cv = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, __fadd_rn((float)py, 0.5f)), p.heightF), 1.0f);
    float d[3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
        d[a] = __fadd_rn(__fadd_rn(__fmul_rn(p.U[a], cu), __fmul_rn(p.V[a], cv)), p.W[a]);
    const float len2 = __fadd_rn(__fadd_rn(__fmul_rn(d[0], d[0]), __fmul_rn(d[1], d[1])), __fmul_rn(d[2], d[2]));
    const float rs = __fdiv_rn(1.0f, __fsqrt_rn(len2));
    float dir[3], t1[3], t2[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        dir[a] = __fmul_rn(d[a], rs);
        const float inv = __fdiv_rn(1.0f, dir[a]);
        t1[a] = __fmul_rn(__fsub_rn(0.f, p.eye[a]), inv);
        t2[a] = __fmul_rn(__fsub_rn(p.bmax[a], p.eye[a]), inv);
    }
    const float tnear = refMax(refMin(t1[0], t2[0]), refMax(refMin(t1[1], t2[1]), refMin(t1[2], t2[2])));
    const float tfar = refMin(refMax(t1[0], t2[0]), refMin(refMax(t1[1], t2[1]), refMax(t1[2], t2[2])));

    float t = tnear;
    float tc[3], inc[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        tc[a] = __fdiv_rn(__fadd_rn(p.eye[a], __fmul_rn(dir[a], t)), p.bmax[a]);
        inc[a] = __fdiv_rn(__fmul_rn(dir[a], p.dt), p.bmax[a]);
    }
    float dst0 = 0.f, dst1 = 0.f, dst2 = 0.f, dst3 = 0.f;
    const bool unitStep = p.dt == 1.0f;                    // pow(x, 1) == x exactly, on both sides
    while (t < tfar) {
        const int ix = (int)__fmul_rn(refClamp(tc[0], 0.0f, p.cX), p.nX);
        const int iy = (int)__fmul_rn(refClamp(tc[1], 0.0f, p.cY), p.nY);
        const int iz = (int)__fmul_rn(refClamp(tc[2], 0.0f, p.cZ), p.nZ);
        const size_t idx = ((size_t)iz * (size_t)p.dimY + (size_t)iy) * (size_t)p.dimX + (size_t)ix;
        const float voxel = normalizeVoxel<F>(fetchVoxel<F>(p, idx));
        float c0, c1, c2, c3;
        if (LUT) {
            const int i = (int)__fmul_rn(refClamp(voxel, 0.0f, p.lutC), p.lutNf);
            const float4 q = __ldg(p.lut + i);
            c0 = q.x; c1 = q.y; c2 = q.z; c3 = q.w;
        } else {
            c0 = c1 = c2 = c3 = voxel;
        }
        const float om = __fsub_rn(1.f, c3);
        c3 = __fsub_rn(1.f, unitStep ? om : powf(om, p.dt));     // opacity correction
        c0 = __fmul_rn(c0, c3); c1 = __fmul_rn(c1, c3); c2 = __fmul_rn(c2, c3);
        const float k = __fsub_rn(1.0f, dst3);
        dst0 = __fadd_rn(dst0, __fmul_rn(c0, k));
        dst1 = __fadd_rn(dst1, __fmul_rn(c1, k));
        dst2 = __fadd_rn(dst2, __fmul_rn(c2, k));
        dst3 = __fadd_rn(dst3, __fmul_rn(c3, k));
#pragma unroll
        for (int a = 0; a < 3; ++a) tc[a] = __fadd_rn(tc[a], inc[a]);
        t = __fadd_rn(t, p.dt);
    }
    if (p.srgb) { dst0 = toSrgb(dst0); dst1 = toSrgb(dst1); dst2 = toSrgb(dst2); }
    const size_t pix = (size_t)py * (size_t)p.width + (size_t)px;
    if (p.rgba32f) p.rgba32f[pix] = make_float4(dst0, dst1, dst2, dst3);
    if (p.rgba8) {
        // float -> unorm<8>: saturate, x 255 in binary64, truncate (visionaray math/detail/unorm.inl:20-24)
        const float o[4] = { dst0, dst1, dst2, dst3 };
        uint32_t w = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float sat = refMax(0.0f, refMin(o[c], 1.0f));
            w |= ((uint32_t)(__dmul_rn((double)sat, 255.0)) & 0xffu) << (8 * c);
        }
        reinterpret_cast<uint32_t*>(p.rgba8)[pix] = w;
    }
}

// host twins of the vector helpers (binary32, one rounding per operation: the file is compiled with -ffp-contract=off)
struct H3 { float x, y, z; };
H3 hsub(H3 a, H3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
H3 hscale(H3 a, float s) { return { a.x * s, a.y * s, a.z * s }; }
H3 hcross(H3 u, H3 v) { return { u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x }; }
float hdot(H3 u, H3 v) { return u.x * v.x + u.y * v.y + u.z * v.z; }
H3 hnormalize(H3 v) { return hscale(v, 1.0f / sqrtf(hdot(v, v))); }

template <int F>
cudaError_t launchF(const LaunchCtx& c, const RenderParams& p)
{
    dim3 grid((unsigned)((p.width + 31) / 32), (unsigned)((p.height + 7) / 8), 1);
    if (p.lut) rayMarchKernel<F, true><<<grid, 256, 0, c.stream>>>(p);
    else rayMarchKernel<F, false><<<grid, 256, 0, c.stream>>>(p);
    countLaunch();
    return cudaGetLastError();
}

} // namespace

bool renderFormatSupported(int format) { return format == F_U8 || format == F_U16 || format == F_INT16 || format == F_U32; }

// pinhole_camera::view_all of the viewer's default camera (Render.cpp:856-865)
void renderDefaultCamera(const int32_t dims[3], const float dist[3], float eye[3], float center[3], float up[3])
{
    const float fovy = 45.f * (float)(3.14159265358979323846 / 180.0);
    const H3 size = { (float)dims[0] * dist[0], (float)dims[1] * dist[1], (float)dims[2] * dist[2] };
    const float r = sqrtf(hdot(size, size)) * 0.5f;
    center[0] = size.x * 0.5f; center[1] = size.y * 0.5f; center[2] = size.z * 0.5f;
    eye[0] = center[0]; eye[1] = center[1]; eye[2] = center[2] + (r + r / atanf(fovy));
    up[0] = 0.f; up[1] = 1.f; up[2] = 0.f;
}

cudaError_t launchRayMarch(const LaunchCtx& c, const VolDesc& vol, const float dist[3], const float eye[3],
                           const float center[3], const float up[3], float fovyDegrees, const float* lutDev, int lutN,
                           float dt, int srgb, int width, int height, uint8_t* rgba8, float* rgba32f)
{
    if (!renderFormatSupported(vol.format) || width <= 0 || height <= 0 || !(dt > 0.f)) return cudaErrorInvalidValue;
    RenderParams p;
    p.vol = vol.data; p.dimX = vol.dims[0]; p.dimY = vol.dims[1]; p.dimZ = vol.dims[2]; p.format = vol.format;
    p.nX = (float)vol.dims[0]; p.nY = (float)vol.dims[1]; p.nZ = (float)vol.dims[2];
    p.cX = 1.0f - 1.0f / p.nX; p.cY = 1.0f - 1.0f / p.nY; p.cZ = 1.0f - 1.0f / p.nZ;
    // pinhole_camera::perspective + begin_frame (visionaray detail/pinhole_camera.inl:31-38, 83-94)
    const float fovy = fovyDegrees * (float)(3.14159265358979323846 / 180.0);
    const float aspect = (float)width / (float)height;
    const H3 e = { eye[0], eye[1], eye[2] }, ce = { center[0], center[1], center[2] }, u0 = { up[0], up[1], up[2] };
    const H3 f = hnormalize(hsub(e, ce));
    const H3 s = hnormalize(hcross(u0, f));
    const H3 u = hcross(f, s);
    const H3 U = hscale(hscale(s, tanf(fovy / 2.0f)), aspect), V = hscale(u, tanf(fovy / 2.0f));
    p.eye[0] = e.x; p.eye[1] = e.y; p.eye[2] = e.z;
    p.U[0] = U.x; p.U[1] = U.y; p.U[2] = U.z;
    p.V[0] = V.x; p.V[1] = V.y; p.V[2] = V.z;
    p.W[0] = -f.x; p.W[1] = -f.y; p.W[2] = -f.z;
    for (int a = 0; a < 3; ++a) p.bmax[a] = (float)vol.dims[a] * dist[a];
    p.lut = (lutDev && lutN > 0) ? reinterpret_cast<const float4*>(lutDev) : nullptr;
    p.lutN = lutN; p.lutNf = (float)lutN; p.lutC = lutN > 0 ? 1.0f - 1.0f / (float)lutN : 0.f;
    p.dt = dt; p.srgb = srgb;
    p.width = width; p.height = height; p.widthF = (float)width; p.heightF = (float)height;
    p.rgba8 = rgba8; p.rgba32f = reinterpret_cast<float4*>(rgba32f);
    switch (vol.format) {
    case F_U8:    return launchF<F_U8>(c, p);
    case F_U16:   return launchF<F_U16>(c, p);
    case F_INT16: return launchF<F_INT16>(c, p);
    case F_U32:   return launchF<F_U32>(c, p);
    }
    return cudaErrorInvalidValue;
}

} // namespace vktb
