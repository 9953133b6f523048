// B200Backend.cpp -- the translation unit a volkit maintainer adds to bind libvktb200.so.
//
// It is compiled HERE against the reference's own, unmodified headers and linked with the reference's own L3 front
// files built with VKT_HAVE_CUDA=1 (integration/Makefile), so the boundary is proven by a compiler and a linker, not
// by prose: every `*_cuda` symbol those fronts reference is defined below as a thin call into the C-ABI
// (include/vktb/backend.h).  It replaces, in src/vkt/CMakeLists.txt:200-217 of the reference:
//     Arithmetic_cuda.cu  Copy_cuda.cu  Fill_cuda.cu  Memory_cuda.cu  Resample_cuda.cu (SV->SV part)  CudaTimer.cpp
// Plain C++14: no nvcc and no CUDA runtime on the volkit side.
#include <chrono>
#include <cstdint>
#include <cstdlib>

#include <vktb/backend.h>

#include <vkt/HierarchicalVolume.hpp>
#include <vkt/Resample.hpp>
#include <vkt/StructuredVolume.hpp>
#include <vkt/Transform.hpp>

#include "CudaTimer.hpp"
#include "HierarchicalVolumeView.hpp"
#include "Logging.hpp"
#include "StructuredVolumeView.hpp"

namespace vkt
{
    // StructuredVolume -> POD descriptor; getData() migrates the volume to the GPU first
    // (include/cpp/vkt/ManagedBuffer.hpp:168-198), exactly as StructuredVolumeView's ctor does today
    static vktb_vol describe(StructuredVolume& v)
    {
        Vec3i d = v.getDims(); Vec2f m = v.getVoxelMapping();
        return vktb_vol{ v.getData(), { d.x, d.y, d.z }, (int32_t)v.getDataFormat(), m.x, m.y };
    }
    static vktb_vol describe(StructuredVolumeView& v)    // FillRange_cuda receives a view (Fill.cpp:60-72)
    {
        Vec3i d = v.getDims(); Vec2f m = v.getVoxelMapping();
        return vktb_vol{ const_cast<uint8_t*>(v.getData()), { d.x, d.y, d.z }, (int32_t)v.getDataFormat(), m.x, m.y };
    }
    static void check(int rc, char const* what)
    {
        if (rc != 0) VKT_LOG(logging::Level::Error) << what << ": " << vktb_last_error();
    }
    static void outOfScope(char const* what)
    {
        VKT_LOG(logging::Level::Error) << what << ": not part of the B200 backend (keep the reference's file for it)";
        std::abort();
    }

    // ---- Memory_cuda.cu ----------------------------------------------------------------------------------------
    void MemsetRange_cuda(void* dst, void const* src, std::size_t dstSize, std::size_t srcSize)
    {
        check(vktb_memset_range(dst, src, dstSize, srcSize), "MemsetRange_cuda");
    }

    // ---- Fill_cuda.cu ------------------------------------------------------------------------------------------
    void FillRange_cuda(StructuredVolumeView volume, Vec3i first, Vec3i last, float value)
    {
        vktb_vol v = describe(volume);
        check(vktb_fill_range(&v, &first.x, &last.x, value), "FillRange_cuda");
    }
    void FillRange_cuda(HierarchicalVolumeView, Vec3i, Vec3i, float) { outOfScope("FillRange_cuda(HierarchicalVolume)"); }

    // ---- Copy_cuda.cu ------------------------------------------------------------------------------------------
    void CopyRange_cuda(StructuredVolume& dst, StructuredVolume& src, Vec3i first, Vec3i last, Vec3i dstOffset)
    {
        vktb_vol d = describe(dst), s = describe(src);
        check(vktb_copy_range(&d, &s, &first.x, &last.x, &dstOffset.x), "CopyRange_cuda");
    }

    // ---- Arithmetic_cuda.cu ------------------------------------------------------------------------------------
#define VKT_B200_ARITH(NAME, OP)                                                                              \
    void NAME##Range_cuda(StructuredVolume& dest, StructuredVolume& source1, StructuredVolume& source2,      \
                          Vec3i first, Vec3i last, Vec3i dstOffset)                                          \
    {                                                                                                        \
        vktb_vol d = describe(dest), a = describe(source1), b = describe(source2);                          \
        check(vktb_arith_range(OP, &d, &a, &b, &first.x, &last.x, &dstOffset.x), #NAME "Range_cuda");       \
    }
    VKT_B200_ARITH(Sum, VKTB_SUM)           VKT_B200_ARITH(Diff, VKTB_DIFF)         VKT_B200_ARITH(Prod, VKTB_PROD)
    VKT_B200_ARITH(Quot, VKTB_QUOT)         VKT_B200_ARITH(AbsDiff, VKTB_ABSDIFF)   VKT_B200_ARITH(SafeSum, VKTB_SAFESUM)
    VKT_B200_ARITH(SafeDiff, VKTB_SAFEDIFF) VKT_B200_ARITH(SafeProd, VKTB_SAFEPROD)
    VKT_B200_ARITH(SafeQuot, VKTB_SAFEQUOT) VKT_B200_ARITH(SafeAbsDiff, VKTB_SAFEABSDIFF)
#undef VKT_B200_ARITH

    // ---- Transform_cuda.hpp (its inline empty stubs become declarations: integration/reference.patch) -----------
    void TransformRange_cuda(StructuredVolume& volume, Vec3i first, Vec3i last, TransformUnaryOp unaryOp)
    {
        int op;   // a host function pointer cannot run on the device: it must have been bound to a device op
        if (vktb_lookup_unary_op((void const*)unaryOp, &op) != 0) { check(-1, "TransformRange_cuda"); return; }
        vktb_vol v = describe(volume);
        check(vktb_transform_range1(&v, &first.x, &last.x, op), "TransformRange_cuda");
    }
    void TransformRange_cuda(StructuredVolume& volume1, StructuredVolume& volume2, Vec3i first, Vec3i last,
                             TransformBinaryOp binaryOp)
    {
        int op;
        if (vktb_lookup_binary_op((void const*)binaryOp, &op) != 0) { check(-1, "TransformRange_cuda"); return; }
        vktb_vol a = describe(volume1), b = describe(volume2);
        check(vktb_transform_range2(&a, &b, &first.x, &last.x, op), "TransformRange_cuda");
    }

    // ---- Resample_cuda.cu --------------------------------------------------------------------------------------
    void Resample_cuda(StructuredVolume& dst, StructuredVolume& src, FilterMode fm)
    {
        vktb_vol d = describe(dst), s = describe(src);
        check(vktb_resample(&d, &s, fm == FilterMode::Linear ? VKTB_LINEAR : VKTB_NEAREST), "Resample_cuda");
    }
    void Resample_cuda(StructuredVolume&, HierarchicalVolume&, FilterMode) { outOfScope("Resample_cuda(HierarchicalVolume)"); }
    void ResampleCLAHE_cuda(StructuredVolume&, StructuredVolume&) { outOfScope("ResampleCLAHE_cuda"); }

    // ---- CudaTimer.cpp -----------------------------------------------------------------------------------------
    // The reference brackets an algorithm with events on the legacy default stream (printPerformance).  The backend
    // queues its kernels on its own stream, so the timer is a host clock around vktb_sync(); the cudaEvent_t slot of
    // the unchanged class layout (CudaTimer.hpp:20-21) carries the start time.
    typedef std::chrono::steady_clock::time_point TimePoint;
    CudaTimer::CudaTimer() : start_(nullptr), stop_(nullptr) { reset(); }
    CudaTimer::~CudaTimer() { delete reinterpret_cast<TimePoint*>(start_); }
    void CudaTimer::reset()
    {
        delete reinterpret_cast<TimePoint*>(start_);
        vktb_sync();
        start_ = reinterpret_cast<cudaEvent_t>(new TimePoint(std::chrono::steady_clock::now()));
    }
    double CudaTimer::getElapsedSeconds() const
    {
        vktb_sync();
        return std::chrono::duration<double>(std::chrono::steady_clock::now() - *reinterpret_cast<TimePoint*>(start_)).count();
    }
} // vkt
