// Front functions of the hot-path algorithms: argument normalisation + dispatch on the calling thread's
// ExecutionPolicy (the reference's VKT_LEGACY_CALL__, src/vkt/Callable.hpp:82-113):
//     Device::CPU + HostAPI::Serial  -> host::xxx (serial.cpp)
//     Device::GPU + DeviceAPI::CUDA  -> the C-ABI (include/vktb/backend.h); a failing call returns InvalidValue
//     anything else                  -> nothing runs (as in the reference), NoError
// With printPerformance the call is timed and one line is logged in the reference's format (Callable.hpp:23-47).
#include <chrono>
#include <cstdio>
#include <vector>

#include <vkt/Aggregates.hpp>
#include <vkt/Arithmetic.hpp>
#include <vkt/Copy.hpp>
#include <vkt/Decompose.hpp>
#include <vkt/Fill.hpp>
#include <vkt/Filter.hpp>
#include <vkt/Flip.hpp>
#include <vkt/Histogram.hpp>
#include <vkt/Resample.hpp>
#include <vkt/Rotate.hpp>
#include <vkt/Scale.hpp>
#include <vkt/Scan.hpp>
#include <vkt/StructuredVolume.hpp>
#include <vkt/Transform.hpp>

#include "backend_link.hpp"
#include "host_codec.hpp"
#include "serial.hpp"

namespace vkt
{
namespace
{
    host::View viewOf(StructuredVolume& v)
    {
        host::View w;
        w.data = v.getData();   // migrates to the host under the CPU policy
        w.dims = v.getDims();
        w.format = v.getDataFormat();
        Vec2f m = v.getVoxelMapping();
        w.lo = m.x; w.hi = m.y;
        w.bpv = host::bytesPerVoxel(w.format);
        return w;
    }

    // Runs cpu() or gpu() according to the policy; gpu() returns a C-ABI status.
    template <class Cpu, class Gpu>
    Error dispatch(char const* name, Cpu cpu, Gpu gpu)
    {
        ExecutionPolicy ep = GetThreadExecutionPolicy();
        typedef std::chrono::steady_clock Clock;
        if (host::onCpuSerial(ep)) {
            Clock::time_point t0 = Clock::now();
            cpu();
            if (ep.printPerformance)
                std::printf("Device: CPU (serial), algorithm: %s_serial, time elapsed: %g sec.\n", name,
                            std::chrono::duration<double>(Clock::now() - t0).count());
            return NoError;
        }
        if (host::onGpu(ep)) {
            if (ep.printPerformance) vktb_sync();
            Clock::time_point t0 = Clock::now();
            if (gpu() != 0) return host::backendError(name);
            if (ep.printPerformance) {
                vktb_sync();
                std::printf("Device: GPU (CUDA), algorithm: %s_cuda, time elapsed: %g sec.\n", name,
                            std::chrono::duration<double>(Clock::now() - t0).count());
            }
            return NoError;
        }
        return NoError;
    }

    void toArray(Vec3i v, int32_t out[3]) { out[0] = v.x; out[1] = v.y; out[2] = v.z; }

    // host streaming (SetHostStreaming): taken when every volume of the call still lives on the host
    bool streamed(StructuredVolume& a) { return GetHostStreaming() && a.residesOnHost(); }
    bool streamed(StructuredVolume& a, StructuredVolume& b) { return streamed(a) && b.residesOnHost(); }
    bool streamed(StructuredVolume& a, StructuredVolume& b, StructuredVolume& c) { return streamed(a, b) && c.residesOnHost(); }
} // namespace

    //---------------------------------------------------------------------------------------------------------
    // Fill (reference src/vkt/Fill.cpp:32-83)
    //---------------------------------------------------------------------------------------------------------
    Error FillRange(StructuredVolume& volume, Vec3i first, Vec3i last, float value)
    {
        return dispatch("FillRange",
            [&] { host::fillRange(viewOf(volume), first, last, value); },
            [&] {
                vktb_vol v = host::describe(volume);
                int32_t f[3], l[3];
                toArray(first, f); toArray(last, l);
                return vktb_fill_range(&v, f, l, value);
            });
    }

    Error FillRange(StructuredVolume& volume, int32_t firstX, int32_t firstY, int32_t firstZ, int32_t lastX,
                    int32_t lastY, int32_t lastZ, float value)
    {
        return FillRange(volume, Vec3i{ firstX, firstY, firstZ }, Vec3i{ lastX, lastY, lastZ }, value);
    }

    Error Fill(StructuredVolume& volume, float value)
    {
        return FillRange(volume, Vec3i{ 0, 0, 0 }, volume.getDims(), value);
    }

    //---------------------------------------------------------------------------------------------------------
    // Copy (reference src/vkt/Copy.cpp:26-95)
    //---------------------------------------------------------------------------------------------------------
    Error CopyRange(StructuredVolume& dst, StructuredVolume& src, Vec3i first, Vec3i last, Vec3i dstOffset)
    {
        return dispatch("CopyRange",
            [&] { host::copyRange(viewOf(dst), viewOf(src), first, last, dstOffset); },
            [&] {
                vktb_vol d = host::describe(dst), s = host::describe(src);
                int32_t f[3], l[3], o[3];
                toArray(first, f); toArray(last, l); toArray(dstOffset, o);
                return vktb_copy_range(&d, &s, f, l, o);
            });
    }

    Error CopyRange(StructuredVolume& dst, StructuredVolume& src, int32_t firstX, int32_t firstY, int32_t firstZ,
                    int32_t lastX, int32_t lastY, int32_t lastZ, int32_t dstOffsetX, int32_t dstOffsetY, int32_t dstOffsetZ)
    {
        return CopyRange(dst, src, Vec3i{ firstX, firstY, firstZ }, Vec3i{ lastX, lastY, lastZ },
                         Vec3i{ dstOffsetX, dstOffsetY, dstOffsetZ });
    }

    Error Copy(StructuredVolume& dst, StructuredVolume& src)
    {
        return CopyRange(dst, src, Vec3i{ 0, 0, 0 }, dst.getDims(), Vec3i{ 0, 0, 0 });   // dst dims: Copy.cpp:32
    }

    //---------------------------------------------------------------------------------------------------------
    // Arithmetic (reference src/vkt/Arithmetic.cpp:26-714); op ids = enum vktb_arith_op
    //---------------------------------------------------------------------------------------------------------
    static Error arithmetic(char const* name, int op, StructuredVolume& dest, StructuredVolume& source1,
                            StructuredVolume& source2, Vec3i first, Vec3i last, Vec3i dstOffset)
    {
        return dispatch(name,
            [&] { host::arithRange(op, viewOf(dest), viewOf(source1), viewOf(source2), first, last, dstOffset); },
            [&] {
                int32_t f[3], l[3], o[3];
                toArray(first, f); toArray(last, l); toArray(dstOffset, o);
                if (streamed(dest, source1, source2)) {
                    vktb_vol d = host::describeInPlace(dest), a = host::describeInPlace(source1),
                             b = host::describeInPlace(source2);
                    return vktb_hosted_arith_range(op, &d, &a, &b, f, l, o);
                }
                vktb_vol d = host::describe(dest), a = host::describe(source1), b = host::describe(source2);
                return vktb_arith_range(op, &d, &a, &b, f, l, o);
            });
    }

#define VKT_DEFINE_ARITHMETIC_OP__(NAME, OP)                                                                    \
    Error NAME##Range(StructuredVolume& dest, StructuredVolume& source1, StructuredVolume& source2, Vec3i first, \
                      Vec3i last, Vec3i dstOffset)                                                              \
    {                                                                                                           \
        return arithmetic(#NAME "Range", OP, dest, source1, source2, first, last, dstOffset);                   \
    }                                                                                                           \
    Error NAME##Range(StructuredVolume& dest, StructuredVolume& source1, StructuredVolume& source2,             \
                      int32_t firstX, int32_t firstY, int32_t firstZ, int32_t lastX, int32_t lastY,             \
                      int32_t lastZ, int32_t dstOffsetX, int32_t dstOffsetY, int32_t dstOffsetZ)                \
    {                                                                                                           \
        return arithmetic(#NAME "Range", OP, dest, source1, source2, Vec3i{ firstX, firstY, firstZ },           \
                          Vec3i{ lastX, lastY, lastZ }, Vec3i{ dstOffsetX, dstOffsetY, dstOffsetZ });           \
    }                                                                                                           \
    Error NAME(StructuredVolume& dest, StructuredVolume& source1, StructuredVolume& source2)                    \
    {                                                                                                           \
        return arithmetic(#NAME "Range", OP, dest, source1, source2, Vec3i{ 0, 0, 0 }, dest.getDims(),          \
                          Vec3i{ 0, 0, 0 });                                                                    \
    }

    VKT_DEFINE_ARITHMETIC_OP__(Sum, VKTB_SUM)
    VKT_DEFINE_ARITHMETIC_OP__(Diff, VKTB_DIFF)
    VKT_DEFINE_ARITHMETIC_OP__(Prod, VKTB_PROD)
    VKT_DEFINE_ARITHMETIC_OP__(Quot, VKTB_QUOT)
    VKT_DEFINE_ARITHMETIC_OP__(AbsDiff, VKTB_ABSDIFF)
    VKT_DEFINE_ARITHMETIC_OP__(SafeSum, VKTB_SAFESUM)
    VKT_DEFINE_ARITHMETIC_OP__(SafeDiff, VKTB_SAFEDIFF)
    VKT_DEFINE_ARITHMETIC_OP__(SafeProd, VKTB_SAFEPROD)
    VKT_DEFINE_ARITHMETIC_OP__(SafeQuot, VKTB_SAFEQUOT)
    VKT_DEFINE_ARITHMETIC_OP__(SafeAbsDiff, VKTB_SAFEABSDIFF)
#undef VKT_DEFINE_ARITHMETIC_OP__

    //---------------------------------------------------------------------------------------------------------
    // Transform (reference src/vkt/Transform.cpp:26-71)
    //---------------------------------------------------------------------------------------------------------
    Error TransformRange(StructuredVolume& volume, Vec3i first, Vec3i last, TransformUnaryOp unaryOp)
    {
        if (!unaryOp) return InvalidValue;
        return dispatch("TransformRange",
            [&] { host::transformRange(viewOf(volume), first, last, unaryOp); },
            [&] {
                int op = -1;
                if (vktb_lookup_unary_op((void const*)unaryOp, &op) != 0) return -1;   // unbound: no CPU fallback
                int32_t f[3], l[3];
                toArray(first, f); toArray(last, l);
                if (streamed(volume)) {
                    vktb_vol v = host::describeInPlace(volume);
                    return vktb_hosted_transform_range1(&v, f, l, op);
                }
                vktb_vol v = host::describe(volume);
                return vktb_transform_range1(&v, f, l, op);
            });
    }

    Error TransformRange(StructuredVolume& volume, int32_t firstX, int32_t firstY, int32_t firstZ, int32_t lastX,
                         int32_t lastY, int32_t lastZ, TransformUnaryOp unaryOp)
    {
        return TransformRange(volume, Vec3i{ firstX, firstY, firstZ }, Vec3i{ lastX, lastY, lastZ }, unaryOp);
    }

    Error Transform(StructuredVolume& volume, TransformUnaryOp unaryOp)
    {
        return TransformRange(volume, Vec3i{ 0, 0, 0 }, volume.getDims(), unaryOp);
    }

    Error TransformRange(StructuredVolume& volume1, StructuredVolume& volume2, Vec3i first, Vec3i last,
                         TransformBinaryOp binaryOp)
    {
        if (!binaryOp) return InvalidValue;
        return dispatch("TransformRange",
            [&] { host::transformRange(viewOf(volume1), viewOf(volume2), first, last, binaryOp); },
            [&] {
                int op = -1;
                if (vktb_lookup_binary_op((void const*)binaryOp, &op) != 0) return -1;
                vktb_vol a = host::describe(volume1), b = host::describe(volume2);
                int32_t f[3], l[3];
                toArray(first, f); toArray(last, l);
                return vktb_transform_range2(&a, &b, f, l, op);
            });
    }

    Error TransformRange(StructuredVolume& volume1, StructuredVolume& volume2, int32_t firstX, int32_t firstY,
                         int32_t firstZ, int32_t lastX, int32_t lastY, int32_t lastZ, TransformBinaryOp binaryOp)
    {
        return TransformRange(volume1, volume2, Vec3i{ firstX, firstY, firstZ }, Vec3i{ lastX, lastY, lastZ }, binaryOp);
    }

    Error Transform(StructuredVolume& volume1, StructuredVolume& volume2, TransformBinaryOp binaryOp)
    {
        return TransformRange(volume1, volume2, Vec3i{ 0, 0, 0 }, volume1.getDims(), binaryOp);
    }

    Error BindDeviceOp(TransformUnaryOp hostOp, char const* deviceOpName)
    {
        int op = -1;
        if (vktb_unary_op_by_name(deviceOpName, &op) != 0 || vktb_bind_unary_op((void const*)hostOp, op) != 0)
            return host::backendError("BindDeviceOp");
        return NoError;
    }

    Error BindDeviceOp(TransformBinaryOp hostOp, char const* deviceOpName)
    {
        int op = -1;
        if (vktb_binary_op_by_name(deviceOpName, &op) != 0 || vktb_bind_binary_op((void const*)hostOp, op) != 0)
            return host::backendError("BindDeviceOp");
        return NoError;
    }

    Error BindDeviceOpImage(TransformUnaryOp hostOp, void const* image, std::size_t size, char const* name)
    {
        int op = -1;
        if (vktb_register_unary_op_image(image, size, name, &op) != 0 || vktb_bind_unary_op((void const*)hostOp, op) != 0)
            return host::backendError("BindDeviceOpImage");
        return NoError;
    }

    Error BindDeviceOpImage(TransformBinaryOp hostOp, void const* image, std::size_t size, char const* name)
    {
        int op = -1;
        if (vktb_register_binary_op_image(image, size, name, &op) != 0 || vktb_bind_binary_op((void const*)hostOp, op) != 0)
            return host::backendError("BindDeviceOpImage");
        return NoError;
    }

    Error BindDeviceOpSource(TransformUnaryOp hostOp, char const* cudaSource, char const* name)
    {
        int op = -1;
        if (vktb_register_unary_op_source(cudaSource, name, &op) != 0 || vktb_bind_unary_op((void const*)hostOp, op) != 0)
            return host::backendError("BindDeviceOpSource");
        return NoError;
    }

    Error BindDeviceOpSource(TransformBinaryOp hostOp, char const* cudaSource, char const* name)
    {
        int op = -1;
        if (vktb_register_binary_op_source(cudaSource, name, &op) != 0 || vktb_bind_binary_op((void const*)hostOp, op) != 0)
            return host::backendError("BindDeviceOpSource");
        return NoError;
    }

    //---------------------------------------------------------------------------------------------------------
    // Resample (reference src/vkt/Resample.cpp:22-31)
    //---------------------------------------------------------------------------------------------------------
    Error Resample(StructuredVolume& dst, StructuredVolume& src, FilterMode fm)
    {
        return dispatch("Resample",
            [&] { host::resample(viewOf(dst), viewOf(src), fm == FilterMode::Linear ? 1 : 0); },
            [&] {
                const int filter = fm == FilterMode::Linear ? VKTB_LINEAR : VKTB_NEAREST;
                if (streamed(dst, src)) {
                    vktb_vol d = host::describeInPlace(dst), s = host::describeInPlace(src);
                    return vktb_hosted_resample(&d, &s, filter);
                }
                vktb_vol d = host::describe(dst), s = host::describe(src);
                return vktb_resample(&d, &s, filter);
            });
    }
    //---------------------------------------------------------------------------------------------------------
    // BrickDecompose (reference src/vkt/Decompose.cpp:22-150, Decompose_serial.hpp:15-48)
    //---------------------------------------------------------------------------------------------------------
    Error host::brickDecompose(StructuredVolume* const* bricks, Vec3i nb, StructuredVolume& source, Vec3i brickSize,
                               Vec3i haloNeg, Vec3i haloPos)
    {
        const std::size_t n = nb.x * std::size_t(nb.y) * nb.z;
        const Vec3i sd = source.getDims();
        std::vector<int32_t> firsts(3 * n), lasts(3 * n), offs(3 * n, 0);
        std::size_t i = 0;
        for (int z = 0; z < nb.z; ++z)
            for (int y = 0; y < nb.y; ++y)
                for (int x = 0; x < nb.x; ++x, ++i) {
                    const int idx[3] = { x, y, z }, bs[3] = { brickSize.x, brickSize.y, brickSize.z };
                    const int dim[3] = { sd.x, sd.y, sd.z }, neg[3] = { haloNeg.x, haloNeg.y, haloNeg.z };
                    const int pos[3] = { haloPos.x, haloPos.y, haloPos.z };
                    for (int a = 0; a < 3; ++a) {
                        const int f = idx[a] * bs[a];
                        firsts[3 * i + a] = f - neg[a];
                        lasts[3 * i + a] = (f + bs[a] < dim[a] ? f + bs[a] : dim[a]) + pos[a];
                    }
                }
        auto v3 = [](int32_t const* p) { return Vec3i{ p[0], p[1], p[2] }; };
        return dispatch("BrickDecompose",
            [&] {
                host::View s = viewOf(source);
                for (std::size_t k = 0; k < n; ++k)
                    host::copyRange(viewOf(*bricks[k]), s, v3(&firsts[3 * k]), v3(&lasts[3 * k]), Vec3i{ 0, 0, 0 });
            },
            [&] {
                // every brick and the source on the device first, then one launch for all boxes
                std::vector<vktb_vol> d(n);
                for (std::size_t k = 0; k < n; ++k) d[k] = host::describe(*bricks[k]);
                vktb_vol s = host::describe(source);
                return vktb_copy_range_batch((int)n, d.data(), &s, firsts.data(), lasts.data(), offs.data());
            });
    }

    Error BrickDecompose(Array3D<StructuredVolume>& dest, StructuredVolume& source, Vec3i brickSize, Vec3i haloSizeNeg,
                         Vec3i haloSizePos)
    {
        std::vector<StructuredVolume*> bricks(dest.numElements());
        for (std::size_t i = 0; i < bricks.size(); ++i) bricks[i] = dest.data() + i;
        return host::brickDecompose(bricks.data(), dest.dims(), source, brickSize, haloSizeNeg, haloSizePos);
    }

    Error BrickDecompose(Array3D<StructuredVolume>& dest, StructuredVolume& source, int32_t brickSizeX, int32_t brickSizeY,
                         int32_t brickSizeZ, int32_t haloSizeNegX, int32_t haloSizeNegY, int32_t haloSizeNegZ,
                         int32_t haloSizePosX, int32_t haloSizePosY, int32_t haloSizePosZ)
    {
        return BrickDecompose(dest, source, Vec3i{ brickSizeX, brickSizeY, brickSizeZ },
                              Vec3i{ haloSizeNegX, haloSizeNegY, haloSizeNegZ }, Vec3i{ haloSizePosX, haloSizePosY, haloSizePosZ });
    }

    Error BrickDecomposeResize(Array3D<StructuredVolume>& dest, StructuredVolume& source, Vec3i brickSize,
                               Vec3i haloSizeNeg, Vec3i haloSizePos)
    {
        if (brickSize.x <= 0 || brickSize.y <= 0 || brickSize.z <= 0) return InvalidValue;
        const Vec3i sd = source.getDims(), nb = host::brickCount(sd, brickSize);
        const Vec3f dist = source.getDist();
        const Vec2f m = source.getVoxelMapping();
        dest = Array3D<StructuredVolume>(nb);
        for (int z = 0; z < nb.z; ++z)
            for (int y = 0; y < nb.y; ++y)
                for (int x = 0; x < nb.x; ++x) {
                    const Vec3i d = host::brickDims(sd, brickSize, haloSizeNeg, haloSizePos, Vec3i{ x, y, z });
                    dest[Vec3i{ x, y, z }] = StructuredVolume(d.x, d.y, d.z, source.getDataFormat(), dist.x, dist.y, dist.z,
                                                              m.x, m.y);
                }
        return NoError;
    }

    Error BrickDecomposeResize(Array3D<StructuredVolume>& dest, StructuredVolume& source, int32_t brickSizeX,
                               int32_t brickSizeY, int32_t brickSizeZ, int32_t haloSizeNegX, int32_t haloSizeNegY,
                               int32_t haloSizeNegZ, int32_t haloSizePosX, int32_t haloSizePosY, int32_t haloSizePosZ)
    {
        return BrickDecomposeResize(dest, source, Vec3i{ brickSizeX, brickSizeY, brickSizeZ },
                                    Vec3i{ haloSizeNegX, haloSizeNegY, haloSizeNegZ },
                                    Vec3i{ haloSizePosX, haloSizePosY, haloSizePosZ });
    }

    //---------------------------------------------------------------------------------------------------------
    // Scan (reference src/vkt/Scan.cpp:22-71)
    //---------------------------------------------------------------------------------------------------------
    Error ScanRange(StructuredVolume& dst, StructuredVolume& src, Vec3i first, Vec3i last, Vec3i dstOffset)
    {
        const bool inPlace = &dst == &src;
        const int f[3] = { first.x, first.y, first.z }, l[3] = { last.x, last.y, last.z };
        for (int a = 0; a < 3; ++a)
            if (f[a] == l[a] || (!inPlace && f[a] > l[a])) return InvalidValue;   // the reference would not terminate
        return dispatch("ScanRange",
            [&] {
                host::View d = viewOf(dst);
                if (!inPlace) host::copyRange(d, viewOf(src), first, last, dstOffset);
                host::scanRange(d, Vec3i{ first.x + dstOffset.x, first.y + dstOffset.y, first.z + dstOffset.z },
                                Vec3i{ last.x + dstOffset.x, last.y + dstOffset.y, last.z + dstOffset.z });
            },
            [&] {
                vktb_vol d = host::describe(dst), s = inPlace ? d : host::describe(src);
                int32_t fa[3], la[3], oa[3];
                toArray(first, fa); toArray(last, la); toArray(dstOffset, oa);
                return vktb_scan_range(&d, &s, fa, la, oa);
            });
    }

    Error ScanRange(StructuredVolume& dst, StructuredVolume& src, int32_t firstX, int32_t firstY, int32_t firstZ,
                    int32_t lastX, int32_t lastY, int32_t lastZ, int32_t dstOffsetX, int32_t dstOffsetY, int32_t dstOffsetZ)
    {
        return ScanRange(dst, src, Vec3i{ firstX, firstY, firstZ }, Vec3i{ lastX, lastY, lastZ },
                         Vec3i{ dstOffsetX, dstOffsetY, dstOffsetZ });
    }

    Error Scan(StructuredVolume& dst, StructuredVolume& src)
    {
        return ScanRange(dst, src, Vec3i{ 0, 0, 0 }, dst.getDims(), Vec3i{ 0, 0, 0 });   // dst dims: Scan.cpp:24-31
    }

    //---------------------------------------------------------------------------------------------------------
    // Flip (reference src/vkt/Flip.cpp:26-117)
    //---------------------------------------------------------------------------------------------------------
    Error FlipRange(StructuredVolume& dest, StructuredVolume& source, Vec3i first, Vec3i last, Vec3i dstOffset, Axis axis)
    {
        if (dest.getBytesPerVoxel() != source.getBytesPerVoxel()) return InvalidValue;
        return dispatch("FlipRange",
            [&] { host::flipRange(viewOf(dest), viewOf(source), first, last, dstOffset, (int)axis); },
            [&] {
                vktb_vol d = host::describe(dest), s = host::describe(source);
                int32_t f[3], l[3], o[3];
                toArray(first, f); toArray(last, l); toArray(dstOffset, o);
                return vktb_flip_range(&d, &s, f, l, o, (int)axis);
            });
    }

    Error FlipRange(StructuredVolume& dest, StructuredVolume& source, Vec3i first, Vec3i last, Axis axis)
    {
        return FlipRange(dest, source, first, last, Vec3i{ 0, 0, 0 }, axis);
    }

    Error FlipRange(StructuredVolume& dest, StructuredVolume& source, int32_t firstX, int32_t firstY, int32_t firstZ,
                    int32_t lastX, int32_t lastY, int32_t lastZ, Axis axis)
    {
        return FlipRange(dest, source, Vec3i{ firstX, firstY, firstZ }, Vec3i{ lastX, lastY, lastZ }, Vec3i{ 0, 0, 0 }, axis);
    }

    Error FlipRange(StructuredVolume& dest, StructuredVolume& source, int32_t firstX, int32_t firstY, int32_t firstZ,
                    int32_t lastX, int32_t lastY, int32_t lastZ, int32_t dstOffsetX, int32_t dstOffsetY, int32_t dstOffsetZ,
                    Axis axis)
    {
        return FlipRange(dest, source, Vec3i{ firstX, firstY, firstZ }, Vec3i{ lastX, lastY, lastZ },
                         Vec3i{ dstOffsetX, dstOffsetY, dstOffsetZ }, axis);
    }

    Error Flip(StructuredVolume& dest, StructuredVolume& source, Axis axis)
    {
        return FlipRange(dest, source, Vec3i{ 0, 0, 0 }, dest.getDims(), Vec3i{ 0, 0, 0 }, axis);   // dest dims: Flip.cpp:28-36
    }

    //---------------------------------------------------------------------------------------------------------
    // Scale / Rotate (reference src/vkt/Scale.cpp:21-40, src/vkt/Rotate.cpp:21-42): range = [0, source dims)
    //---------------------------------------------------------------------------------------------------------
    namespace
    {
        Error affine(char const* name, StructuredVolume& dest, StructuredVolume& source, const float m[9], Vec3f center)
        {
            if (&dest == &source) return InvalidValue;      // the serial walk would read its own output
            const Vec3i first{ 0, 0, 0 }, last = source.getDims();
            return dispatch(name,
                [&] { host::affineRange(viewOf(dest), viewOf(source), first, last, m, center); },
                [&] {
                    vktb_vol d = host::describe(dest), s = host::describe(source);
                    int32_t f[3], l[3];
                    toArray(first, f); toArray(last, l);
                    const float c[3] = { center.x, center.y, center.z };
                    return vktb_affine_range(&d, &s, f, l, m, c);
                });
        }
    }

    Error Scale(StructuredVolume& dest, StructuredVolume& source, Vec3f scalingFactor, Vec3f centerOfScaling)
    {
        float m[9];
        host::scaleMatrix(scalingFactor, m);
        return affine("ScaleRange", dest, source, m, centerOfScaling);
    }

    Error Rotate(StructuredVolume& dest, StructuredVolume& source, Vec3f axis, float angleInRadians, Vec3f centerOfRotation)
    {
        float m[9];
        host::rotateMatrix(axis, angleInRadians, m);
        return affine("RotateRange", dest, source, m, centerOfRotation);
    }

    //---------------------------------------------------------------------------------------------------------
    // ApplyFilter (reference src/vkt/Filter.cpp:25-66)
    //---------------------------------------------------------------------------------------------------------
    Filter::Filter(float* data, Vec3i dims)
        : data_(data, data + (dims.x > 0 && dims.y > 0 && dims.z > 0 ? dims.x * std::size_t(dims.y) * dims.z : 0)), dims_(dims) {}
    float* Filter::getData() { return data_.data(); }
    Vec3i Filter::getDims() const { return dims_; }

    Error ApplyFilter(StructuredVolume& dest, StructuredVolume& source, Filter& filter, AddressMode am)
    {
        const Vec3i fd = filter.getDims();
        if (fd.x < 1 || fd.x > 3 || fd.y < 1 || fd.y > 3 || fd.z < 1 || fd.z > 3) return InvalidValue;
        const Vec3i dd = dest.getDims(), sd = source.getDims();
        if (dd.x < sd.x || dd.y < sd.y || dd.z < sd.z) return InvalidValue;
        return dispatch("ApplyFilterRange",
            [&] {
                host::View d = viewOf(dest);
                host::applyFilter(d, &dest == &source ? d : viewOf(source), filter.getData(), fd);
            },
            [&] {
                vktb_vol d = host::describe(dest), s = &dest == &source ? d : host::describe(source);
                int32_t f[3];
                toArray(fd, f);
                return vktb_apply_filter(&d, &s, filter.getData(), f, (int)am);
            });
    }

    //---------------------------------------------------------------------------------------------------------
    // Aggregates (reference src/vkt/Aggregates.cpp:23-66)
    //---------------------------------------------------------------------------------------------------------
    Error ComputeAggregatesRange(StructuredVolume& volume, Aggregates& aggregates, Vec3i first, Vec3i last)
    {
        static_assert(sizeof(Aggregates) == sizeof(vktb_aggregates), "vkt::Aggregates is the C-ABI struct");
        if (first.x > last.x || first.y > last.y || first.z > last.z) return InvalidValue;   // the reference would not end
        return dispatch("ComputeAggregatesRange",
            [&] { host::aggregatesRange(viewOf(volume), first, last, aggregates); },
            [&] {
                vktb_vol v = host::describe(volume);
                int32_t f[3], l[3];
                toArray(first, f); toArray(last, l);
                return vktb_aggregates_range(&v, f, l, reinterpret_cast<vktb_aggregates*>(&aggregates));
            });
    }

    Error ComputeAggregatesRange(StructuredVolume& volume, Aggregates& aggregates, int32_t firstX, int32_t firstY,
                                 int32_t firstZ, int32_t lastX, int32_t lastY, int32_t lastZ)
    {
        return ComputeAggregatesRange(volume, aggregates, Vec3i{ firstX, firstY, firstZ }, Vec3i{ lastX, lastY, lastZ });
    }

    Error ComputeAggregates(StructuredVolume& volume, Aggregates& aggregates)
    {
        return ComputeAggregatesRange(volume, aggregates, Vec3i{ 0, 0, 0 }, volume.getDims());
    }

    //---------------------------------------------------------------------------------------------------------
    // Histogram (reference src/vkt/Histogram.cpp:23-84)
    //---------------------------------------------------------------------------------------------------------
    Histogram::Histogram(std::size_t numBins) { resize(numBins); }
    std::size_t Histogram::getNumBins() const { return size_; }
    std::size_t* Histogram::getBinCounts() { migrate(); return data_; }

    Error ComputeHistogramRange(StructuredVolume& volume, Histogram& histogram, Vec3i first, Vec3i last)
    {
        static_assert(sizeof(std::size_t) == sizeof(uint64_t), "bins are 64-bit counters");
        return dispatch("ComputeHistogramRange",
            [&] {
                host::View v = viewOf(volume);
                host::histogramRange(v, first, last, histogram.getBinCounts(), histogram.getNumBins());
            },
            [&] {
                vktb_vol v = host::describe(volume);
                int32_t f[3], l[3];
                toArray(first, f); toArray(last, l);
                return vktb_histogram_range(&v, f, l, (uint64_t*)histogram.getBinCounts(), histogram.getNumBins(), nullptr);
            });
    }

    Error ComputeHistogramRange(StructuredVolume& volume, Histogram& histogram, int32_t firstX, int32_t firstY,
                                int32_t firstZ, int32_t lastX, int32_t lastY, int32_t lastZ)
    {
        return ComputeHistogramRange(volume, histogram, Vec3i{ firstX, firstY, firstZ }, Vec3i{ lastX, lastY, lastZ });
    }

    Error ComputeHistogram(StructuredVolume& volume, Histogram& histogram)
    {
        return ComputeHistogramRange(volume, histogram, Vec3i{ 0, 0, 0 }, volume.getDims());
    }
} // vkt
