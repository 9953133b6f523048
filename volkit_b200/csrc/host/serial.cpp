#include "serial.hpp"

#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>

#include "host_codec.hpp"

namespace vkt
{
namespace host
{
    static inline int clampi(int x, int a, int b) { int m = b < x ? b : x; return a < m ? m : a; }

    // reference src/vkt/Fill_serial.hpp:20-26: every voxel of [first,last) receives map(value)
    void fillRange(View v, Vec3i first, Vec3i last, float value)
    {
        uint8_t raw[8] = {};
        mapVoxel(raw, value, v.format, v.lo, v.hi);
        for (int32_t z = first.z; z < last.z; ++z)
            for (int32_t y = first.y; y < last.y; ++y) {
                uint8_t* row = v.data + v.offset(first.x, y, z);
                for (int32_t x = first.x; x < last.x; ++x, row += v.bpv) std::memcpy(row, raw, v.bpv);
            }
    }

    // reference src/vkt/Copy_serial.hpp:13-82: dst[p - first + off] = src[clamp(p)], bytes iff format and mapping agree
    void copyRange(View dst, View src, Vec3i first, Vec3i last, Vec3i off)
    {
        const bool bytes = dst.format == src.format && dst.lo == src.lo && dst.hi == src.hi;
        for (int32_t z = first.z; z < last.z; ++z)
            for (int32_t y = first.y; y < last.y; ++y)
                for (int32_t x = first.x; x < last.x; ++x) {
                    const uint8_t* s = src.data + src.offset(clampi(x, 0, src.dims.x - 1), clampi(y, 0, src.dims.y - 1),
                                                             clampi(z, 0, src.dims.z - 1));
                    uint8_t* d = dst.data + dst.offset(x - first.x + off.x, y - first.y + off.y, z - first.z + off.z);
                    if (bytes) std::memcpy(d, s, dst.bpv);
                    else mapVoxel(d, unmapVoxel(s, src.format, src.lo, src.hi), dst.format, dst.lo, dst.hi);
                }
    }

    // reference src/vkt/Arithmetic_serial.hpp:15-45 (loop) and :47-260 (the ten ops; op ids in header order)
    void arithRange(int op, View dst, View s1, View s2, Vec3i first, Vec3i last, Vec3i off)
    {
        for (int32_t z = first.z; z < last.z; ++z)
            for (int32_t y = first.y; y < last.y; ++y)
                for (int32_t x = first.x; x < last.x; ++x) {
                    const float a = unmapVoxel(s1.data + s1.offset(x, y, z), s1.format, s1.lo, s1.hi);
                    const float b = unmapVoxel(s2.data + s2.offset(x, y, z), s2.format, s2.lo, s2.hi);
                    float r;
                    switch (op % 5) {
                    case 0: r = a + b; break;
                    case 1: r = a - b; break;
                    case 2: r = a * b; break;
                    case 3: r = a / b; break;
                    default: r = std::fabs(a - b); break;
                    }
                    if (op >= 5) r = refClamp(r, dst.lo, dst.hi);
                    mapVoxel(dst.data + dst.offset(x + off.x, y + off.y, z + off.z), r, dst.format, dst.lo, dst.hi);
                }
    }

    // reference src/vkt/Transform_serial.hpp:15-50
    void transformRange(View v, Vec3i first, Vec3i last, TransformUnaryOp op)
    {
        for (int32_t z = first.z; z < last.z; ++z)
            for (int32_t y = first.y; y < last.y; ++y)
                for (int32_t x = first.x; x < last.x; ++x) {
                    uint8_t scratch[8] = {};
                    uint8_t* p = v.data + v.offset(x, y, z);
                    std::memcpy(scratch, p, v.bpv);
                    op(x, y, z, VoxelView{ scratch, v.format, v.lo, v.hi });
                    std::memcpy(p, scratch, v.bpv);
                }
    }

    // reference src/vkt/Transform_serial.hpp:53-103
    void transformRange(View v1, View v2, Vec3i first, Vec3i last, TransformBinaryOp op)
    {
        for (int32_t z = first.z; z < last.z; ++z)
            for (int32_t y = first.y; y < last.y; ++y)
                for (int32_t x = first.x; x < last.x; ++x) {
                    uint8_t s1[8] = {}, s2[8] = {};
                    uint8_t* p1 = v1.data + v1.offset(x, y, z);
                    uint8_t* p2 = v2.data + v2.offset(x, y, z);
                    std::memcpy(s1, p1, v1.bpv);
                    std::memcpy(s2, p2, v2.bpv);
                    op(x, y, z, VoxelView{ s1, v1.format, v1.lo, v1.hi }, VoxelView{ s2, v2.format, v2.lo, v2.hi });
                    std::memcpy(p1, s1, v1.bpv);
                    std::memcpy(p2, s2, v2.bpv);
                }
    }

    // One tap of the source as a float; linear indices outside the buffer (only reachable through the reference's
    // unclamped +1 x tap at the very last voxel) read as a zero voxel.
    static inline float tap(const View& s, long long index)
    {
        const long long n = (long long)s.dims.x * s.dims.y * s.dims.z;
        uint8_t zero[8] = {};
        const uint8_t* p = (index >= 0 && index < n) ? s.data + (std::size_t)index * s.bpv : zero;
        return unmapVoxel(p, s.format, s.lo, s.hi);
    }

    // reference src/vkt/Resample_serial.hpp:26-71 and StructuredVolumeView::sampleLinear
    // (src/vkt/StructuredVolumeView.hpp:80-119): filter 0 = Nearest, 1 = Linear as the reference computes it (integer
    // parameters -> fraction 0; lo taps effectively unclamped, hi.y / hi.z clamped, hi.x not)
    void resample(View dst, View src, int filter)
    {
        const Vec3i D = dst.dims, S = src.dims;
        if (D.x == S.x && D.y == S.y && D.z == S.z) {
            const std::size_t n = (std::size_t)D.x * D.y * D.z;
            for (std::size_t i = 0; i < n; ++i)
                mapVoxel(dst.data + i * dst.bpv, unmapVoxel(src.data + i * src.bpv, src.format, src.lo, src.hi),
                         dst.format, dst.lo, dst.hi);
            return;
        }
        const long long row = S.x, plane = (long long)S.x * S.y;
        for (int32_t z = 0; z < D.z; ++z) {
            const int sz = cvtt32((float)z / (float)D.z * (float)S.z);
            for (int32_t y = 0; y < D.y; ++y) {
                const int sy = cvtt32((float)y / (float)D.y * (float)S.y);
                for (int32_t x = 0; x < D.x; ++x) {
                    const int sx = cvtt32((float)x / (float)D.x * (float)S.x);
                    float value;
                    if (filter == 0) {
                        value = tap(src, sz * plane + sy * row + sx);
                    } else {
                        // lo.x / lo.y clamped, lo.z not; hi.y / hi.z clamped, hi.x not (the reference's typo'd clamps)
                        const int lx = clampi(sx, 0, S.x - 1), ly = clampi(sy, 0, S.y - 1), lz = sz;
                        const int hx = sx + 1, hy = clampi(sy + 1, 0, S.y - 1), hz = clampi(sz + 1, 0, S.z - 1);
                        const float fx = (float)sx - (float)lx, fy = (float)sy - (float)ly, fz = (float)sz - (float)lz;
                        const float v0 = tap(src, lz * plane + ly * row + lx), v1 = tap(src, lz * plane + ly * row + hx);
                        const float v2 = tap(src, lz * plane + hy * row + lx), v3 = tap(src, lz * plane + hy * row + hx);
                        const float v4 = tap(src, hz * plane + ly * row + lx), v5 = tap(src, hz * plane + ly * row + hx);
                        const float v6 = tap(src, hz * plane + hy * row + lx), v7 = tap(src, hz * plane + hy * row + hx);
                        value = refLerp(refLerp(refLerp(v0, v1, fx), refLerp(v2, v3, fx), fy),
                                        refLerp(refLerp(v4, v5, fx), refLerp(v6, v7, fx), fy), fz);
                    }
                    mapVoxel(dst.data + dst.offset(x, y, z), value, dst.format, dst.lo, dst.hi);
                }
            }
        }
    }
    // StructuredVolumeView::sampleLinear(int32_t, int32_t, int32_t) (src/vkt/StructuredVolumeView.hpp:80-119), clamps as written
    static float sampleLinearRef(const View& src, int sx, int sy, int sz)
    {
        const Vec3i S = src.dims;
        const long long row = S.x, plane = (long long)S.x * S.y;
        const int lx = clampi(sx, 0, S.x - 1), ly = clampi(sy, 0, S.y - 1), lz = sz;
        const int hx = sx + 1, hy = clampi(sy + 1, 0, S.y - 1), hz = clampi(sz + 1, 0, S.z - 1);
        const float fx = (float)sx - (float)lx, fy = (float)sy - (float)ly, fz = (float)sz - (float)lz;
        const float v0 = tap(src, lz * plane + ly * row + lx), v1 = tap(src, lz * plane + ly * row + hx);
        const float v2 = tap(src, lz * plane + hy * row + lx), v3 = tap(src, lz * plane + hy * row + hx);
        const float v4 = tap(src, hz * plane + ly * row + lx), v5 = tap(src, hz * plane + ly * row + hx);
        const float v6 = tap(src, hz * plane + hy * row + lx), v7 = tap(src, hz * plane + hy * row + hx);
        return refLerp(refLerp(refLerp(v0, v1, fx), refLerp(v2, v3, fx), fy),
                       refLerp(refLerp(v4, v5, fx), refLerp(v6, v7, fx), fy), fz);
    }

    // reference src/vkt/Scale_serial.hpp:27-31
    void scaleMatrix(Vec3f f, float m[9])
    {
        for (int i = 0; i < 9; ++i) m[i] = 0.f;
        m[0] = 1.f / f.x; m[4] = 1.f / f.y; m[8] = 1.f / f.z;
    }

    // reference src/vkt/Rotate_serial.hpp:55-72 (normalised axis, inverted angle, quaternion) and :17-39 (to a matrix)
    void rotateMatrix(Vec3f axis, float angle, float m[9])
    {
        const float len = sqrtf(axis.x * axis.x + axis.y * axis.y + axis.z * axis.z);
        const float ax = axis.x / len, ay = axis.y / len, az = axis.z / len;
        angle = 6.28319f - angle;
        const float re = cosf(angle * .5f);
        const float ix = ax * sinf(angle * .5f), iy = ay * sinf(angle * .5f), iz = az * sinf(angle * .5f);
        const float xx = ix * ix, xy = ix * iy, xz = ix * iz, xw = ix * re, yy = iy * iy, yz = iy * iz, yw = iy * re;
        const float zz = iz * iz, zw = iz * re, ww = re * re;
        m[0] = 2.f * (ww + xx) - 1.f; m[1] = 2.f * (xy + zw);       m[2] = 2.f * (xz - yw);
        m[3] = 2.f * (xy - zw);       m[4] = 2.f * (ww + yy) - 1.f; m[5] = 2.f * (yz + xw);
        m[6] = 2.f * (xz + yw);       m[7] = 2.f * (yz - xw);       m[8] = 2.f * (ww + zz) - 1.f;
    }

    // reference src/vkt/Scale_serial.hpp:36-55 / Rotate_serial.hpp:77-93: the whole dst volume is walked
    void affineRange(View dst, View src, Vec3i first, Vec3i last, const float m[9], Vec3f c)
    {
        for (int32_t z = 0; z < dst.dims.z; ++z)
            for (int32_t y = 0; y < dst.dims.y; ++y)
                for (int32_t x = 0; x < dst.dims.x; ++x) {
                    const float px = (float)x - c.x, py = (float)y - c.y, pz = (float)z - c.z;
                    float qx = m[0] * px + m[3] * py + m[6] * pz;
                    float qy = m[1] * px + m[4] * py + m[7] * pz;
                    float qz = m[2] * px + m[5] * py + m[8] * pz;
                    qx = qx + c.x; qy = qy + c.y; qz = qz + c.z;
                    if (qx >= (float)first.x && qx < (float)last.x && qy >= (float)first.y && qy < (float)last.y
                        && qz >= (float)first.z && qz < (float)last.z)
                        mapVoxel(dst.data + dst.offset(x, y, z), sampleLinearRef(src, cvtt32(qx), cvtt32(qy), cvtt32(qz)),
                                 dst.format, dst.lo, dst.hi);
                }
    }

    // reference src/vkt/Filter_serial.hpp:15-64 over [0, src dims)
    void applyFilter(View dst, View src, const float* w, Vec3i fd)
    {
        const std::size_t bytes = (std::size_t)src.dims.x * src.dims.y * src.dims.z * src.bpv;
        std::vector<uint8_t> copy(src.data, src.data + bytes);
        View temp = src;
        temp.data = copy.data();
        const Vec3i last = src.dims;
        const int hx = fd.x / 2, hy = fd.y / 2, hz = fd.z / 2;
        for (int32_t z = 0; z < last.z; ++z)
            for (int32_t y = 0; y < last.y; ++y)
                for (int32_t x = 0; x < last.x; ++x) {
                    float value = 0.f;
                    if (!(x == 0 || y == 0 || z == 0 || x == last.x - 1 || y == last.y - 1 || z == last.z - 1))
                        for (int32_t zz = 0; zz < fd.z; ++zz)
                            for (int32_t yy = 0; yy < fd.y; ++yy)
                                for (int32_t xx = 0; xx < fd.x; ++xx)
                                    value += w[(zz * fd.y + yy) * fd.x + xx]
                                             * unmapVoxel(temp.data + temp.offset(x - hx + xx, y - hy + yy, z - hz + zz),
                                                          temp.format, temp.lo, temp.hi);
                    mapVoxel(dst.data + dst.offset(x, y, z), value, dst.format, dst.lo, dst.hi);
                }
    }

    // reference src/vkt/Flip_serial.hpp:11-63, as written (first ignored, z mirrored on Axis::Y, Axis::Z mirrors nothing)
    void flipRange(View dst, View src, Vec3i first, Vec3i last, Vec3i off, int axis)
    {
        int32_t rx = last.x - first.x, ry = last.y - first.y, rz = last.z - first.z;
        if (axis == 0) rx >>= 1; else if (axis == 1) ry >>= 1; else rz >>= 1;
        for (int32_t z = 0; z < rz; ++z)
            for (int32_t y = 0; y < ry; ++y)
                for (int32_t x = 0; x < rx; ++x) {
                    const int32_t xx = axis == 0 ? last.x - 1 - x : x;
                    const int32_t yy = axis == 1 ? last.y - 1 - y : y;
                    const int32_t zz = axis == 1 ? last.z - 1 - z : z;
                    uint8_t v1[8], v2[8];
                    std::memcpy(v1, src.data + src.offset(x, y, z), src.bpv);
                    std::memcpy(v2, src.data + src.offset(xx, yy, zz), src.bpv);
                    std::memcpy(dst.data + dst.offset(off.x + x, off.y + y, off.z + z), v2, dst.bpv);
                    std::memcpy(dst.data + dst.offset(off.x + xx, off.y + yy, off.z + zz), v1, dst.bpv);
                }
    }

    // reference src/vkt/Aggregates_serial.hpp:20-83: binary32 running accumulators, x fastest; the divisions by the
    // WHOLE volume's voxel count (:60) run in binary64 and are stored back as floats
    void aggregatesRange(View vol, Vec3i first, Vec3i last, Aggregates& a)
    {
        std::memset(&a, 0, sizeof(a));
        a.min = FLT_MAX;
        a.max = -FLT_MAX;
        a.prod = 1.f;
        for (int32_t z = first.z; z < last.z; ++z)
            for (int32_t y = first.y; y < last.y; ++y)
                for (int32_t x = first.x; x < last.x; ++x) {
                    const float val = unmapVoxel(vol.data + vol.offset(x, y, z), vol.format, vol.lo, vol.hi);
                    if (val < a.min) { a.min = val; a.argmin = { x, y, z }; }
                    if (val > a.max) { a.max = val; a.argmax = { x, y, z }; }
                    a.mean += val;
                    a.sum += val;
                    a.prod *= val;
                }
        const std::size_t numElems = vol.dims.x * std::size_t(vol.dims.y) * vol.dims.z;
        a.mean = (float)(a.mean / (double)numElems);
        for (int32_t z = first.z; z < last.z; ++z)
            for (int32_t y = first.y; y < last.y; ++y)
                for (int32_t x = first.x; x < last.x; ++x) {
                    const float val = unmapVoxel(vol.data + vol.offset(x, y, z), vol.format, vol.lo, vol.hi);
                    a.var += (val - a.mean) * (val - a.mean);
                }
        a.var = (float)(a.var / (double)numElems);
        a.stddev = sqrtf(a.var);
    }

    // reference src/vkt/Histogram_serial.hpp:18-51; indices outside [0, numBins) are skipped (the reference's UB)
    void histogramRange(View vol, Vec3i first, Vec3i last, std::size_t* bins, std::size_t numBins)
    {
        for (std::size_t i = 0; i < numBins; ++i) bins[i] = 0;
        const float scale = (float)numBins / (vol.hi - vol.lo);
        for (int32_t z = first.z; z < last.z; ++z)
            for (int32_t y = first.y; y < last.y; ++y)
                for (int32_t x = first.x; x < last.x; ++x) {
                    const float val = unmapVoxel(vol.data + vol.offset(x, y, z), vol.format, vol.lo, vol.hi);
                    const float b = (val - vol.lo) * scale;
                    if (b >= 0.0f && b < (float)numBins) bins[(std::size_t)b]++;
                    else if (b > -1.0f && b < 0.0f) bins[0]++;
                }
    }

    //--- Scan: reference src/vkt/Scan_serial.hpp:18-152 ------------------------------------------------------------
    // The reference's inclusion-exclusion over raw voxel integers (64-bit sums, low bpv bytes stored) equals three
    // 1-D inclusive prefix sums modulo 2^(8 bpv): along x, then y, then z, each from `first` towards `last`.
    void scanRange(View v, Vec3i first, Vec3i last)
    {
        const int f[3] = { first.x, first.y, first.z }, l[3] = { last.x, last.y, last.z };
        int n[3], step[3];
        for (int a = 0; a < 3; ++a) {
            step[a] = f[a] < l[a] ? 1 : -1;
            n[a] = step[a] > 0 ? l[a] - f[a] : f[a] - l[a];
        }
        const long long pitch[3] = { (long long)v.bpv * step[0], (long long)v.bpv * v.dims.x * step[1],
                                     (long long)v.bpv * v.dims.x * v.dims.y * step[2] };
        uint8_t* corner = v.data + v.offset(f[0], f[1], f[2]);
        for (int axis = 0; axis < 3; ++axis) {
            const int a1 = (axis + 1) % 3, a2 = (axis + 2) % 3;
            for (int j = 0; j < n[a2]; ++j)
                for (int i = 0; i < n[a1]; ++i) {
                    uint8_t* e = corner + i * pitch[a1] + j * pitch[a2];
                    uint64_t acc = 0;
                    for (int s = 0; s < n[axis]; ++s, e += pitch[axis]) {
                        uint64_t raw = 0;
                        std::memcpy(&raw, e, (std::size_t)v.bpv);
                        acc += raw;
                        std::memcpy(e, &acc, (std::size_t)v.bpv);
                    }
                }
        }
    }

    //--- BrickDecomposeResize layout: reference src/vkt/Decompose.cpp:96-146 ------------------------------------
    Vec3i brickCount(Vec3i d, Vec3i b)
    {
        return Vec3i{ (d.x + b.x - 1) / b.x, (d.y + b.y - 1) / b.y, (d.z + b.z - 1) / b.z };
    }

    Vec3i brickDims(Vec3i d, Vec3i b, Vec3i neg, Vec3i pos, Vec3i index)
    {
        Vec3i n = brickCount(d, b);
        // brick size (without halos) at the far borders: what is left of the volume there
        Vec3i border{ d.x % b.x == 0 ? b.x : b.x - n.x * b.x + d.x,
                      d.y % b.y == 0 ? b.y : b.y - n.y * b.y + d.y,
                      d.z % b.z == 0 ? b.z : b.z - n.z * b.z + d.z };
        return Vec3i{ neg.x + (index.x < n.x - 1 ? b.x : border.x) + pos.x,
                      neg.y + (index.y < n.y - 1 ? b.y : border.y) + pos.y,
                      neg.z + (index.z < n.z - 1 ? b.z : border.z) + pos.z };
    }
} // host
} // vkt
