// C99 API over the C++ host mirror (reference: the "C API" halves of src/vkt/{ExecutionPolicy,Memory,
// StructuredVolume,Voxel,Fill,Copy,Arithmetic,Transform}.cpp).  The opaque handle wraps a vkt::StructuredVolume.
#include <vkt/Aggregates.h>
#include <vkt/Arithmetic.h>
#include <vkt/Array3D.h>
#include <vkt/Copy.h>
#include <vkt/Decompose.h>
#include <vkt/ExecutionPolicy.h>
#include <vkt/Fill.h>
#include <vkt/Flip.h>
#include <vkt/Memory.h>
#include <vkt/Resample.h>
#include <vkt/Rotate.h>
#include <vkt/Scale.h>
#include <vkt/Scan.h>
#include <vkt/StructuredVolume.h>
#include <vkt/Transform.h>
#include <vkt/Voxel.h>

#include <vkt/Aggregates.hpp>
#include <vkt/Arithmetic.hpp>
#include <vkt/Copy.hpp>
#include <vkt/Decompose.hpp>
#include <vkt/ExecutionPolicy.hpp>
#include <vkt/Fill.hpp>
#include <vkt/Flip.hpp>
#include <vkt/Memory.hpp>
#include <vkt/Resample.hpp>
#include <vkt/Rotate.hpp>
#include <vkt/Scale.hpp>
#include <vkt/Scan.hpp>
#include <vkt/StructuredVolume.hpp>
#include <vkt/Transform.hpp>
#include <vkt/Voxel.hpp>

#include <cstring>
#include <vector>

#include "serial.hpp"

struct vktStructuredVolume_impl
{
    vktStructuredVolume_impl(int32_t x, int32_t y, int32_t z, vkt::DataFormat f, float dx, float dy, float dz, float lo, float hi)
        : volume(x, y, z, f, dx, dy, dz, lo, hi) {}
    explicit vktStructuredVolume_impl(vkt::StructuredVolume& rhs) : volume(rhs) {}
    vkt::StructuredVolume volume;
};

struct vktArray3D_vktStructuredVolume_impl
{
    std::vector<vktStructuredVolume> slots;   // x-fastest; NULL until a volume is created in the slot
    vktVec3i_t dims;
};

static size_t slotCount(vktVec3i_t d) { return d.x * (size_t)d.y * d.z; }

// vktVoxelView_t and vkt::VoxelView are layout-identical PODs passed by value: the op pointer is re-typed, as the
// reference does (src/vkt/Transform.cpp:79-153)
static vkt::TransformUnaryOp unaryCast(vktTransformUnaryOp op) { return reinterpret_cast<vkt::TransformUnaryOp>(reinterpret_cast<void (*)(void)>(op)); }
static vkt::TransformBinaryOp binaryCast(vktTransformBinaryOp op) { return reinterpret_cast<vkt::TransformBinaryOp>(reinterpret_cast<void (*)(void)>(op)); }

extern "C" {

//--- ExecutionPolicy ------------------------------------------------------------------------------------------
void vktSetThreadExecutionPolicy(vktExecutionPolicy_t policy)
{
    vkt::ExecutionPolicy ep;
    ep.device = (vkt::ExecutionPolicy::Device)policy.device;
    ep.hostApi = (vkt::ExecutionPolicy::HostAPI)policy.hostApi;
    ep.deviceApi = (vkt::ExecutionPolicy::DeviceAPI)policy.deviceApi;
    ep.printPerformance = policy.printPerformance;
    vkt::SetThreadExecutionPolicy(ep);
}

void vktSetHostStreaming(vktBool_t enabled) { vkt::SetHostStreaming(enabled != 0); }
vktBool_t vktGetHostStreaming(void) { return vkt::GetHostStreaming() ? 1 : 0; }

vktExecutionPolicy_t vktGetThreadExecutionPolicy(void)
{
    vkt::ExecutionPolicy ep = vkt::GetThreadExecutionPolicy();
    vktExecutionPolicy_t p;
    p.device = (vktExecutionPolicyDevice)ep.device;
    p.hostApi = (vktExecutionPolicyHostAPI)ep.hostApi;
    p.deviceApi = (vktExecutionPolicyDeviceAPI)ep.deviceApi;
    p.printPerformance = ep.printPerformance;
    return p;
}

//--- Memory ----------------------------------------------------------------------------------------------------
void vktAllocate(void** ptr, size_t size) { vkt::Allocate(ptr, size); }
void vktFree(void* ptr) { vkt::Free(ptr); }
void vktMemcpy(void* dst, void const* src, size_t size, vktCopyKind ck) { vkt::Memcpy(dst, src, size, (vkt::CopyKind)ck); }
void vktMemsetRange(void* dst, void const* src, size_t dstSize, size_t srcSize) { vkt::MemsetRange(dst, src, dstSize, srcSize); }

//--- Voxel -----------------------------------------------------------------------------------------------------
vktError vktMapVoxel(uint8_t* dst, float value, vktDataFormat f, float lo, float hi)
{
    return (vktError)vkt::MapVoxel(dst, value, (vkt::DataFormat)f, lo, hi);
}

vktError vktUnmapVoxel(float* value, uint8_t const* src, vktDataFormat f, float lo, float hi)
{
    return (vktError)vkt::UnmapVoxel(*value, src, (vkt::DataFormat)f, lo, hi);
}

//--- StructuredVolume ------------------------------------------------------------------------------------------
uint8_t vktStructuredVolumeGetMaxBytesPerVoxel(void) { return vkt::StructuredVolume::GetMaxBytesPerVoxel(); }

void vktStructuredVolumeCreate(vktStructuredVolume* volume, int32_t dimX, int32_t dimY, int32_t dimZ, vktDataFormat f,
                               float distX, float distY, float distZ, float lo, float hi)
{
    *volume = new vktStructuredVolume_impl(dimX, dimY, dimZ, (vkt::DataFormat)f, distX, distY, distZ, lo, hi);
}

void vktStructuredVolumeCreateCopy(vktStructuredVolume* volume, vktStructuredVolume rhs) { *volume = new vktStructuredVolume_impl(rhs->volume); }
void vktStructuredVolumeDestroy(vktStructuredVolume volume) { delete volume; }

void vktStructuredVolumeSetDims3i(vktStructuredVolume v, int32_t x, int32_t y, int32_t z) { v->volume.setDims(x, y, z); }
void vktStructuredVolumeGetDims3i(vktStructuredVolume v, int32_t* x, int32_t* y, int32_t* z) { v->volume.getDims(*x, *y, *z); }
void vktStructuredVolumeSetDims3iv(vktStructuredVolume v, vktVec3i_t d) { v->volume.setDims(vkt::Vec3i{ d.x, d.y, d.z }); }
vktVec3i_t vktStructuredVolumeGetDims3iv(vktStructuredVolume v) { vkt::Vec3i d = v->volume.getDims(); return vktVec3i_t{ d.x, d.y, d.z }; }
void vktStructuredVolumeSetDataFormat(vktStructuredVolume v, vktDataFormat f) { v->volume.setDataFormat((vkt::DataFormat)f); }
vktDataFormat vktStructuredVolumeGetDataFormat(vktStructuredVolume v) { return (vktDataFormat)v->volume.getDataFormat(); }
void vktStructuredVolumeSetDist3f(vktStructuredVolume v, float x, float y, float z) { v->volume.setDist(x, y, z); }
void vktStructuredVolumeGetDist3f(vktStructuredVolume v, float* x, float* y, float* z) { v->volume.getDist(*x, *y, *z); }
void vktStructuredVolumeSetDist3fv(vktStructuredVolume v, vktVec3f_t d) { v->volume.setDist(vkt::Vec3f{ d.x, d.y, d.z }); }
vktVec3f_t vktStructuredVolumeGetDist3fv(vktStructuredVolume v) { vkt::Vec3f d = v->volume.getDist(); return vktVec3f_t{ d.x, d.y, d.z }; }
void vktStructuredVolumeSetVoxelMapping2f(vktStructuredVolume v, float lo, float hi) { v->volume.setVoxelMapping(lo, hi); }
void vktStructuredVolumeGetVoxelMapping2f(vktStructuredVolume v, float* lo, float* hi) { v->volume.getVoxelMapping(*lo, *hi); }
void vktStructuredVolumeSetVoxelMapping2fv(vktStructuredVolume v, vktVec2f_t m) { v->volume.setVoxelMapping(vkt::Vec2f{ m.x, m.y }); }
vktVec2f_t vktStructuredVolumeGetVoxelMapping2fv(vktStructuredVolume v) { vkt::Vec2f m = v->volume.getVoxelMapping(); return vktVec2f_t{ m.x, m.y }; }

static vktBox3f_t toC(vkt::Box3f b) { return vktBox3f_t{ { b.min.x, b.min.y, b.min.z }, { b.max.x, b.max.y, b.max.z } }; }
vktBox3f_t vktStructuredVolumeGetDomainBounds(vktStructuredVolume v) { return toC(v->volume.getDomainBounds()); }
vktBox3f_t vktStructuredVolumeGetObjectBounds(vktStructuredVolume v) { return toC(v->volume.getObjectBounds()); }

uint8_t* vktStructuredVolumeGetData(vktStructuredVolume v) { return v->volume.getData(); }
void vktStructuredVolumeSetValue(vktStructuredVolume v, int32_t x, int32_t y, int32_t z, float value) { v->volume.setValue(x, y, z, value); }
void vktStructuredVolumeGetValue(vktStructuredVolume v, int32_t x, int32_t y, int32_t z, float* value) { v->volume.getValue(x, y, z, *value); }
void vktStructuredVolumeSetBytes(vktStructuredVolume v, int32_t x, int32_t y, int32_t z, uint8_t const* d) { v->volume.setBytes(x, y, z, d); }
void vktStructuredVolumeGetBytes(vktStructuredVolume v, int32_t x, int32_t y, int32_t z, uint8_t* d) { v->volume.getBytes(x, y, z, d); }
uint8_t vktStructuredVolumeGetBytesPerVoxel(vktStructuredVolume v) { return v->volume.getBytesPerVoxel(); }
size_t vktStructuredVolumeGetSizeInBytes(vktStructuredVolume v) { return v->volume.getSizeInBytes(); }
vktResourceHandle vktStructuredVolumeGetResourceHandle(vktStructuredVolume v) { return v->volume.getResourceHandle(); }
void vktStructuredVolumeMigrate(vktStructuredVolume v) { v->volume.migrate(); }

//--- Fill / Copy -----------------------------------------------------------------------------------------------
vktError vktFillSV(vktStructuredVolume v, float value) { return (vktError)vkt::Fill(v->volume, value); }
vktError vktFillRangeSV(vktStructuredVolume v, int32_t fx, int32_t fy, int32_t fz, int32_t lx, int32_t ly, int32_t lz, float value)
{
    return (vktError)vkt::FillRange(v->volume, fx, fy, fz, lx, ly, lz, value);
}

vktError vktCopySV(vktStructuredVolume dst, vktStructuredVolume src) { return (vktError)vkt::Copy(dst->volume, src->volume); }
vktError vktCopyRangeSV(vktStructuredVolume dst, vktStructuredVolume src, int32_t fx, int32_t fy, int32_t fz, int32_t lx,
                        int32_t ly, int32_t lz, int32_t ox, int32_t oy, int32_t oz)
{
    return (vktError)vkt::CopyRange(dst->volume, src->volume, fx, fy, fz, lx, ly, lz, ox, oy, oz);
}

//--- Array3D of volume handles (reference include/c/vkt/Array3D.h:18-231) -------------------------------------
void vktArray3D_vktStructuredVolume_CreateEmpty(vktArray3D_vktStructuredVolume* arr)
{
    *arr = new vktArray3D_vktStructuredVolume_impl();
    (*arr)->dims = vktVec3i_t{ 0, 0, 0 };
}
void vktArray3D_vktStructuredVolume_Create(vktArray3D_vktStructuredVolume* arr, vktVec3i_t dims)
{
    *arr = new vktArray3D_vktStructuredVolume_impl();
    (*arr)->slots.assign(slotCount(dims), nullptr);
    (*arr)->dims = dims;
}
void vktArray3D_vktStructuredVolume_CreateCopy(vktArray3D_vktStructuredVolume* arr, vktArray3D_vktStructuredVolume rhs)
{
    *arr = new vktArray3D_vktStructuredVolume_impl(*rhs);   // copies the handles, like the reference's byte copy
}
void vktArray3D_vktStructuredVolume_Destroy(vktArray3D_vktStructuredVolume arr) { delete arr; }
void vktArray3D_vktStructuredVolume_Resize(vktArray3D_vktStructuredVolume arr, vktVec3i_t dims)
{
    arr->slots.resize(slotCount(dims), nullptr);
    arr->dims = dims;
}
void vktArray3D_vktStructuredVolume_Fill(vktArray3D_vktStructuredVolume arr, vktStructuredVolume value)
{
    arr->slots.assign(arr->slots.size(), value);
}
vktStructuredVolume* vktArray3D_vktStructuredVolume_Data(vktArray3D_vktStructuredVolume arr) { return arr->slots.data(); }
vktStructuredVolume const* vktArray3D_vktStructuredVolume_CData(vktArray3D_vktStructuredVolume arr) { return arr->slots.data(); }
vktStructuredVolume* vktArray3D_vktStructuredVolume_Begin(vktArray3D_vktStructuredVolume arr) { return arr->slots.data(); }
vktStructuredVolume const* vktArray3D_vktStructuredVolume_CBegin(vktArray3D_vktStructuredVolume arr) { return arr->slots.data(); }
vktStructuredVolume* vktArray3D_vktStructuredVolume_End(vktArray3D_vktStructuredVolume arr) { return arr->slots.data() + arr->slots.size(); }
vktStructuredVolume const* vktArray3D_vktStructuredVolume_CEnd(vktArray3D_vktStructuredVolume arr) { return arr->slots.data() + arr->slots.size(); }
vktStructuredVolume* vktArray3D_vktStructuredVolume_Access(vktArray3D_vktStructuredVolume arr, vktVec3i_t i)
{
    return arr->slots.data() + (i.z * arr->dims.x * (size_t)arr->dims.y + i.y * (size_t)arr->dims.x + i.x);
}
vktStructuredVolume const* vktArray3D_vktStructuredVolume_CAccess(vktArray3D_vktStructuredVolume arr, vktVec3i_t i)
{
    return vktArray3D_vktStructuredVolume_Access(arr, i);
}
vktBool_t vktArray3D_vktStructuredVolume_Empty(vktArray3D_vktStructuredVolume arr) { return arr->slots.empty() ? 1 : 0; }
vktVec3i_t vktArray3D_vktStructuredVolume_Dims(vktArray3D_vktStructuredVolume arr) { return arr->dims; }
size_t vktArray3D_vktStructuredVolume_NumElements(vktArray3D_vktStructuredVolume arr) { return arr->slots.size(); }

//--- BrickDecompose (reference src/vkt/Decompose.cpp:152-256) ---------------------------------------------------
vktError vktBrickDecomposeSV(vktArray3D_vktStructuredVolume decomp, vktStructuredVolume source, int32_t bx, int32_t by,
                             int32_t bz, int32_t nx, int32_t ny, int32_t nz, int32_t px, int32_t py, int32_t pz)
{
    std::vector<vkt::StructuredVolume*> bricks(decomp->slots.size());
    for (size_t i = 0; i < bricks.size(); ++i) {
        if (!decomp->slots[i]) return vktInvalidValue;
        bricks[i] = &decomp->slots[i]->volume;
    }
    vkt::Vec3i nb{ decomp->dims.x, decomp->dims.y, decomp->dims.z };
    return (vktError)vkt::host::brickDecompose(bricks.data(), nb, source->volume, vkt::Vec3i{ bx, by, bz },
                                               vkt::Vec3i{ nx, ny, nz }, vkt::Vec3i{ px, py, pz });
}

vktError vktBrickDecomposeResizeSV(vktArray3D_vktStructuredVolume decomp, vktStructuredVolume source, int32_t bx,
                                   int32_t by, int32_t bz, int32_t nx, int32_t ny, int32_t nz, int32_t px, int32_t py,
                                   int32_t pz)
{
    if (bx <= 0 || by <= 0 || bz <= 0) return vktInvalidValue;
    vkt::StructuredVolume& s = source->volume;
    vkt::Vec3i sd = s.getDims(), bs{ bx, by, bz }, nb = vkt::host::brickCount(sd, bs);
    vktArray3D_vktStructuredVolume_Resize(decomp, vktVec3i_t{ nb.x, nb.y, nb.z });
    vkt::Vec3f dist = s.getDist();
    vkt::Vec2f m = s.getVoxelMapping();
    for (int z = 0; z < nb.z; ++z)
        for (int y = 0; y < nb.y; ++y)
            for (int x = 0; x < nb.x; ++x) {
                vkt::Vec3i d = vkt::host::brickDims(sd, bs, vkt::Vec3i{ nx, ny, nz }, vkt::Vec3i{ px, py, pz }, vkt::Vec3i{ x, y, z });
                vktStructuredVolumeCreate(vktArray3D_vktStructuredVolume_Access(decomp, vktVec3i_t{ x, y, z }), d.x, d.y, d.z,
                                          (vktDataFormat)s.getDataFormat(), dist.x, dist.y, dist.z, m.x, m.y);
            }
    return vktNoError;
}

//--- Scale / Rotate (reference src/vkt/Scale.cpp:47-65, src/vkt/Rotate.cpp:49-69) ------------------------------------
vktError vktScaleSV(vktStructuredVolume dest, vktStructuredVolume source, vktVec3f_t scalingFactor, vktVec3f_t centerOfScaling)
{
    return (vktError)vkt::Scale(dest->volume, source->volume, { scalingFactor.x, scalingFactor.y, scalingFactor.z },
                                { centerOfScaling.x, centerOfScaling.y, centerOfScaling.z });
}
vktError vktRotateSV(vktStructuredVolume dest, vktStructuredVolume source, vktVec3f_t axis, float angleInRadians,
                     vktVec3f_t centerOfRotation)
{
    return (vktError)vkt::Rotate(dest->volume, source->volume, { axis.x, axis.y, axis.z }, angleInRadians,
                                 { centerOfRotation.x, centerOfRotation.y, centerOfRotation.z });
}

//--- Aggregates (reference src/vkt/Aggregates.cpp:75-112) ---------------------------------------------------------
vktError vktComputeAggregatesSV(vktStructuredVolume volume, vktAggregates_t* aggregates)
{
    static_assert(sizeof(vktAggregates_t) == sizeof(vkt::Aggregates), "same 13 words");
    vkt::Aggregates a;
    vktError e = (vktError)vkt::ComputeAggregates(volume->volume, a);
    std::memcpy(aggregates, &a, sizeof(a));
    return e;
}
vktError vktComputeAggregatesRangeSV(vktStructuredVolume volume, vktAggregates_t* aggregates, int32_t fx, int32_t fy,
                                     int32_t fz, int32_t lx, int32_t ly, int32_t lz)
{
    vkt::Aggregates a;
    vktError e = (vktError)vkt::ComputeAggregatesRange(volume->volume, a, fx, fy, fz, lx, ly, lz);
    std::memcpy(aggregates, &a, sizeof(a));
    return e;
}

//--- Scan (reference src/vkt/Scan.cpp:79-122) -------------------------------------------------------------------
vktError vktScanSV(vktStructuredVolume dst, vktStructuredVolume src)
{
    return (vktError)vkt::Scan(dst->volume, src->volume);
}
vktError vktScanRangeSV(vktStructuredVolume dst, vktStructuredVolume src, int32_t fx, int32_t fy, int32_t fz, int32_t lx,
                        int32_t ly, int32_t lz, int32_t ox, int32_t oy, int32_t oz)
{
    return (vktError)vkt::ScanRange(dst->volume, src->volume, fx, fy, fz, lx, ly, lz, ox, oy, oz);
}

//--- Flip (reference src/vkt/Flip.cpp:127-167) -------------------------------------------------------------------
vktError vktFlipSV(vktStructuredVolume dest, vktStructuredVolume source, vktAxis axis)
{
    return (vktError)vkt::Flip(dest->volume, source->volume, (vkt::Axis)axis);
}
vktError vktFlipRangeSV(vktStructuredVolume dest, vktStructuredVolume source, int32_t fx, int32_t fy, int32_t fz, int32_t lx,
                        int32_t ly, int32_t lz, int32_t ox, int32_t oy, int32_t oz, vktAxis axis)
{
    return (vktError)vkt::FlipRange(dest->volume, source->volume, fx, fy, fz, lx, ly, lz, ox, oy, oz, (vkt::Axis)axis);
}

//--- Arithmetic ------------------------------------------------------------------------------------------------
#define VKT_DEFINE_ARITHMETIC_OP_C__(NAME)                                                                          \
    vktError vkt##NAME##SV(vktStructuredVolume d, vktStructuredVolume a, vktStructuredVolume b)                     \
    {                                                                                                               \
        return (vktError)vkt::NAME(d->volume, a->volume, b->volume);                                                \
    }                                                                                                               \
    vktError vkt##NAME##RangeSV(vktStructuredVolume d, vktStructuredVolume a, vktStructuredVolume b, int32_t fx,    \
                                int32_t fy, int32_t fz, int32_t lx, int32_t ly, int32_t lz, int32_t ox, int32_t oy, \
                                int32_t oz)                                                                         \
    {                                                                                                               \
        return (vktError)vkt::NAME##Range(d->volume, a->volume, b->volume, fx, fy, fz, lx, ly, lz, ox, oy, oz);     \
    }

VKT_DEFINE_ARITHMETIC_OP_C__(Sum)
VKT_DEFINE_ARITHMETIC_OP_C__(Diff)
VKT_DEFINE_ARITHMETIC_OP_C__(Prod)
VKT_DEFINE_ARITHMETIC_OP_C__(Quot)
VKT_DEFINE_ARITHMETIC_OP_C__(AbsDiff)
VKT_DEFINE_ARITHMETIC_OP_C__(SafeSum)
VKT_DEFINE_ARITHMETIC_OP_C__(SafeDiff)
VKT_DEFINE_ARITHMETIC_OP_C__(SafeProd)
VKT_DEFINE_ARITHMETIC_OP_C__(SafeQuot)
VKT_DEFINE_ARITHMETIC_OP_C__(SafeAbsDiff)
#undef VKT_DEFINE_ARITHMETIC_OP_C__

//--- Transform (vktVoxelView_t and vkt::VoxelView share one layout; the op pointer is cast as in the reference,
//    src/vkt/Transform.cpp:79-153) -----------------------------------------------------------------------------
vktError vktTransformSV1(vktStructuredVolume v, vktTransformUnaryOp op) { return (vktError)vkt::Transform(v->volume, unaryCast(op)); }
vktError vktTransformSV2(vktStructuredVolume v1, vktStructuredVolume v2, vktTransformBinaryOp op)
{
    return (vktError)vkt::Transform(v1->volume, v2->volume, binaryCast(op));
}
vktError vktTransformRangeSV1(vktStructuredVolume v, int32_t fx, int32_t fy, int32_t fz, int32_t lx, int32_t ly, int32_t lz,
                              vktTransformUnaryOp op)
{
    return (vktError)vkt::TransformRange(v->volume, fx, fy, fz, lx, ly, lz, unaryCast(op));
}
vktError vktTransformRangeSV2(vktStructuredVolume v1, vktStructuredVolume v2, int32_t fx, int32_t fy, int32_t fz, int32_t lx,
                              int32_t ly, int32_t lz, vktTransformBinaryOp op)
{
    return (vktError)vkt::TransformRange(v1->volume, v2->volume, fx, fy, fz, lx, ly, lz, binaryCast(op));
}
vktError vktBindDeviceOp1(vktTransformUnaryOp op, char const* name) { return (vktError)vkt::BindDeviceOp(unaryCast(op), name); }
vktError vktBindDeviceOp2(vktTransformBinaryOp op, char const* name) { return (vktError)vkt::BindDeviceOp(binaryCast(op), name); }

//--- Resample --------------------------------------------------------------------------------------------------
vktError vktResampleSV(vktStructuredVolume dst, vktStructuredVolume src, vktFilterMode fm)
{
    return (vktError)vkt::Resample(dst->volume, src->volume, (vkt::FilterMode)fm);
}

} // extern "C"
