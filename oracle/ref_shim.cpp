// TEST INFRASTRUCTURE ONLY -- never linked into, imported by or shipped with the product
// (volkit_b200/).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may load the library this file is built into.
//
// Flat extern "C" driver over the UNMODIFIED reference (Kniggi/volkit) serial CPU path.
// The reference sources are compiled where they lie under /root/reference by oracle/Makefile;
// this file only *calls* the reference's public C++ API (include/cpp/vkt/*.hpp) so that Python
// (ctypes) can run the reference on caller-provided byte buffers:
//
//   vkt::FillRange      include/cpp/vkt/Fill.hpp        -> src/vkt/Fill.cpp:56-83
//   vkt::CopyRange      include/cpp/vkt/Copy.hpp        -> src/vkt/Copy.cpp:26-121
//   vkt::*Range (x10)   include/cpp/vkt/Arithmetic.hpp  -> src/vkt/Arithmetic.cpp:26-714
//   vkt::TransformRange include/cpp/vkt/Transform.hpp   -> src/vkt/Transform.cpp:26-71
//   vkt::Resample       include/cpp/vkt/Resample.hpp    -> src/vkt/Resample.cpp:22-31
//   vkt::MapVoxel/UnmapVoxel include/cpp/vkt/Voxel.hpp  -> src/vkt/Voxel.cpp
//
// Every call runs under the reference's default thread policy {CPU, Serial}.

#include <malloc.h>
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <vector>

#include <vkt/Aggregates.hpp>
#include <vkt/Arithmetic.hpp>
#include <vkt/Copy.hpp>
#include <vkt/Decompose.hpp>
#include <vkt/Decompose.h>
#include <vkt/StructuredVolume.h>
#include <vkt/ExecutionPolicy.hpp>
#include <vkt/Fill.hpp>
#include <vkt/Filter.hpp>
#include <vkt/Flip.hpp>
#include <vkt/Histogram.hpp>
#include <vkt/InputStream.hpp>
#include <vkt/OutputStream.hpp>
#include <vkt/RawFile.hpp>
#include <vkt/Resample.hpp>
#include <vkt/Rotate.hpp>
#include <vkt/Scale.hpp>
#include <vkt/Scan.hpp>
#include <vkt/StructuredVolume.hpp>
#include <vkt/Transform.hpp>
#include <vkt/Voxel.hpp>

#include "StructuredVolume_impl.hpp" // reference-internal: the struct behind the C handle
#include "Transform_serial.hpp" // reference-internal: src/vkt/Transform_serial.hpp (binary overload)

namespace {

// POD volume descriptor shared with the ctypes side (same field order as vktb_vol, include/vktb/backend.h)
struct RefVol
{
    void*   data;
    int32_t dims[3];
    int32_t format;    // vkt::DataFormat numeric value
    float   lo, hi;
};

// A reference volume that holds a private copy of the caller's bytes.
struct Staged
{
    vkt::StructuredVolume vol;
    RefVol const*         desc;
    size_t                bytes;

    explicit Staged(RefVol const* d)
        : vol(d->dims[0], d->dims[1], d->dims[2], (vkt::DataFormat)d->format, 1.f, 1.f, 1.f, d->lo, d->hi)
        , desc(d)
    {
        bytes = vol.getSizeInBytes();
        if (bytes) std::memcpy(vol.getData(), d->data, bytes);
        // The reference's sampleLinear reads ONE voxel past the end of a source volume (its unclamped hi.x tap in the last
        // row of the last plane, weight 0: StructuredVolumeView.hpp:90-99).  That voxel lies in the slack of the malloc
        // chunk (Allocate_serial is a plain malloc), i.e. it holds whatever an earlier allocation left there -- an Inf or
        // NaN bit pattern turns the result of the tapping dst voxels into NaN and made a test flaky.  Zero the slack:
        // the tap then reads a finite 0, which is also what the restatement defines it to be.
        if (bytes) {
            const size_t usable = malloc_usable_size(vol.getData());
            if (usable > bytes) std::memset((char*)vol.getData() + bytes, 0, usable - bytes);
        }
    }

    void writeBack() { if (bytes) std::memcpy(desc->data, vol.getData(), bytes); }
};

typedef vkt::Error (*ArithFn)(vkt::StructuredVolume&, vkt::StructuredVolume&, vkt::StructuredVolume&,
                              vkt::Vec3i, vkt::Vec3i, vkt::Vec3i);

ArithFn arithTable(int op)
{
    switch (op)
    {
    case 0: return (ArithFn)&vkt::SumRange;
    case 1: return (ArithFn)&vkt::DiffRange;
    case 2: return (ArithFn)&vkt::ProdRange;
    case 3: return (ArithFn)&vkt::QuotRange;
    case 4: return (ArithFn)&vkt::AbsDiffRange;
    case 5: return (ArithFn)&vkt::SafeSumRange;
    case 6: return (ArithFn)&vkt::SafeDiffRange;
    case 7: return (ArithFn)&vkt::SafeProdRange;
    case 8: return (ArithFn)&vkt::SafeQuotRange;
    case 9: return (ArithFn)&vkt::SafeAbsDiffRange;
    }
    return nullptr;
}

// ---- Transform ops used by the parity tests (same source text is compiled for the device in
// volkit_b200/csrc/kernels/transform_ops.cuh; the semantics come from SURVEY.md 8(d) and
// src/examples/CoreAlgorithms.c:20-24 / src/examples/Arithmetic.cpp:9-21). -----------------

int bytesPerVoxel(vkt::DataFormat f)
{
    switch (f)
    {
    case vkt::DataFormat::Int8:  case vkt::DataFormat::UInt8:  return 1;
    case vkt::DataFormat::Int16: case vkt::DataFormat::UInt16: return 2;
    default: return 4;
    }
}

// op 0: bytes[0] ^= (x+y+z)
void opXorCoord(int32_t x, int32_t y, int32_t z, vkt::VoxelView v)
{
    v.bytes[0] ^= (uint8_t)(x + y + z);
}

// op 1: bitwise invert every byte of the voxel
void opInvert(int32_t, int32_t, int32_t, vkt::VoxelView v)
{
    int n = bytesPerVoxel(v.dataFormat);
    for (int i = 0; i < n; ++i) v.bytes[i] = (uint8_t)~v.bytes[i];
}

// op 2: 3-D checkerboard of cell size 8: black cells are zeroed, white cells untouched
void opChecker8(int32_t x, int32_t y, int32_t z, vkt::VoxelView v)
{
    if ((((x >> 3) + (y >> 3) + (z >> 3)) & 1) == 0)
    {
        int n = bytesPerVoxel(v.dataFormat);
        for (int i = 0; i < n; ++i) v.bytes[i] = 0;
    }
}

// op 3: threshold through the voxel mapping: value < 0.5*(lo+hi) -> lo else hi
void opThreshold(int32_t, int32_t, int32_t, vkt::VoxelView v)
{
    float val = 0.f;
    vkt::UnmapVoxel(val, v.bytes, v.dataFormat, v.mappingLo, v.mappingHi);
    float mid = 0.5f * (v.mappingLo + v.mappingHi);
    float out = val < mid ? v.mappingLo : v.mappingHi;
    vkt::MapVoxel(v.bytes, out, v.dataFormat, v.mappingLo, v.mappingHi);
}

// binary op 0: v1 |= v2; v2 = v1   (bytewise)
void op2OrAssign(int32_t, int32_t, int32_t, vkt::VoxelView v1, vkt::VoxelView v2)
{
    int n = bytesPerVoxel(v1.dataFormat);
    for (int i = 0; i < n; ++i) { v1.bytes[i] |= v2.bytes[i]; v2.bytes[i] = v1.bytes[i]; }
}

// binary op 1: swap the two voxels bytewise
void op2Swap(int32_t, int32_t, int32_t, vkt::VoxelView v1, vkt::VoxelView v2)
{
    int n = bytesPerVoxel(v1.dataFormat);
    for (int i = 0; i < n; ++i) { uint8_t t = v1.bytes[i]; v1.bytes[i] = v2.bytes[i]; v2.bytes[i] = t; }
}

vkt::TransformUnaryOp unaryTable(int op)
{
    switch (op)
    {
    case 0: return &opXorCoord;
    case 1: return &opInvert;
    case 2: return &opChecker8;
    case 3: return &opThreshold;
    }
    return nullptr;
}

vkt::TransformBinaryOp binaryTable(int op)
{
    switch (op)
    {
    case 0: return &op2OrAssign;
    case 1: return &op2Swap;
    }
    return nullptr;
}

vkt::Vec3i v3(int32_t const* p) { return { p[0], p[1], p[2] }; }

} // namespace

extern "C" {

int ref_abi_version() { return 1; }

void ref_map_voxel(uint8_t* dst, float value, int fmt, float lo, float hi)
{
    vkt::MapVoxel(dst, value, (vkt::DataFormat)fmt, lo, hi);
}

float ref_unmap_voxel(uint8_t const* src, int fmt, float lo, float hi)
{
    float v = 0.f;
    vkt::UnmapVoxel(v, src, (vkt::DataFormat)fmt, lo, hi);
    return v;
}

int ref_fill_range(RefVol const* vol, int32_t const* first, int32_t const* last, float value)
{
    Staged s(vol);
    vkt::FillRange(s.vol, v3(first), v3(last), value);
    s.writeBack();
    return 0;
}

int ref_copy_range(RefVol const* dst, RefVol const* src, int32_t const* first, int32_t const* last,
                   int32_t const* off)
{
    Staged d(dst), s(src);
    vkt::CopyRange(d.vol, s.vol, v3(first), v3(last), v3(off));
    d.writeBack();
    return 0;
}

int ref_arith_range(int op, RefVol const* dst, RefVol const* s1, RefVol const* s2,
                    int32_t const* first, int32_t const* last, int32_t const* off)
{
    ArithFn fn = arithTable(op);
    if (!fn) return -1;
    Staged d(dst), a(s1), b(s2);
    fn(d.vol, a.vol, b.vol, v3(first), v3(last), v3(off));
    d.writeBack();
    return 0;
}

int ref_transform_range1(RefVol const* vol, int32_t const* first, int32_t const* last, int op)
{
    vkt::TransformUnaryOp fn = unaryTable(op);
    if (!fn) return -1;
    Staged s(vol);
    vkt::TransformRange(s.vol, v3(first), v3(last), fn);
    s.writeBack();
    return 0;
}

// A caller-supplied op (tests/test_user_ops.py: the host twin of a VKTB_UNARY_OP, compiled by g++) run by the
// UNMODIFIED reference: vkt::TransformRange with that function pointer.
int ref_transform_fn1(RefVol const* vol, int32_t const* first, int32_t const* last, void* fn)
{
    if (!fn) return -1;
    Staged s(vol);
    vkt::TransformRange(s.vol, v3(first), v3(last), (vkt::TransformUnaryOp)fn);
    s.writeBack();
    return 0;
}

int ref_transform_fn2(RefVol const* vol1, RefVol const* vol2, int32_t const* first, int32_t const* last, void* fn)
{
    if (!fn) return -1;
    Staged a(vol1), b(vol2);
    vkt::TransformRange_serial(a.vol, b.vol, v3(first), v3(last), (vkt::TransformBinaryOp)fn);
    a.writeBack();
    b.writeBack();
    return 0;
}

// The reference declares but never defines the C++ binary overloads (include/cpp/vkt/Transform.hpp:30-62);
// the binary form is only reachable through the C API (src/vkt/Transform.cpp:117-153).
// so the shim calls the serial body directly (src/vkt/Transform_serial.hpp:53-103), which is what
// vktTransformRangeSV2 dispatches to under the CPU policy.
int ref_transform_range2(RefVol const* vol1, RefVol const* vol2, int32_t const* first, int32_t const* last, int op)
{
    vkt::TransformBinaryOp fn = binaryTable(op);
    if (!fn) return -1;
    Staged a(vol1), b(vol2);
    vkt::TransformRange_serial(a.vol, b.vol, v3(first), v3(last), fn);
    a.writeBack();
    b.writeBack();
    return 0;
}

// filter: 0 = Nearest, 1 = Linear (include/cpp/vkt/Resample.hpp:14-18)
int ref_resample(RefVol const* dst, RefVol const* src, int filter)
{
    // Note: under Linear the reference's unclamped hi.x tap reads one voxel past the end of src at
    // the very last src voxel (StructuredVolumeView.hpp:90-99); the value is multiplied by weight 0.
    // The read stays inside the malloc chunk (glibc rounds chunks up to 16n+8 bytes); Staged zeroes that slack.
    Staged d(dst), s(src);
    vkt::Resample(d.vol, s.vol, filter ? vkt::FilterMode::Linear : vkt::FilterMode::Nearest);
    d.writeBack();
    return 0;
}

// vkt::FlipRange (include/cpp/vkt/Flip.hpp -> src/vkt/Flip.cpp:94-104).  Same data pointer for dst and src = the
// in-place call on ONE volume object.
int ref_flip_range(RefVol const* dst, RefVol const* src, int32_t const* first, int32_t const* last,
                   int32_t const* off, int axis)
{
    if (axis < 0 || axis > 2) return -1;
    if (dst->data == src->data) {
        Staged v(dst);
        vkt::FlipRange(v.vol, v.vol, v3(first), v3(last), v3(off), (vkt::Axis)axis);
        v.writeBack();
        return 0;
    }
    Staged d(dst), s(src);
    vkt::FlipRange(d.vol, s.vol, v3(first), v3(last), v3(off), (vkt::Axis)axis);
    d.writeBack();
    return 0;
}

// vkt::ComputeHistogramRange (include/cpp/vkt/Histogram.hpp -> src/vkt/Histogram.cpp:60-91).  The caller must keep
// every voxel's bin inside [0, num_bins): the reference writes out of bounds otherwise.
int ref_histogram_range(RefVol const* vol, int32_t const* first, int32_t const* last, uint64_t* bins, size_t numBins,
                        uint64_t* dropped)
{
    Staged v(vol);
    vkt::Histogram h(numBins);
    vkt::ComputeHistogramRange(v.vol, h, v3(first), v3(last));
    std::size_t* c = h.getBinCounts();
    for (size_t i = 0; i < numBins; ++i) bins[i] = (uint64_t)c[i];
    if (dropped) *dropped = 0;
    return 0;
}

// vkt::Scale (include/cpp/vkt/Scale.hpp:12-15 -> src/vkt/Scale.cpp:21-40), vkt::Rotate (include/cpp/vkt/Rotate.hpp:12-16
// -> src/vkt/Rotate.cpp:21-42), vkt::ApplyFilter (include/cpp/vkt/Filter.hpp:42-45 -> src/vkt/Filter.cpp:48-66)
int ref_scale(RefVol const* dst, RefVol const* src, float const* factor, float const* center)
{
    if (dst->data == src->data) return -1;
    Staged d(dst), s(src);
    vkt::Scale(d.vol, s.vol, { factor[0], factor[1], factor[2] }, { center[0], center[1], center[2] });
    d.writeBack();
    return 0;
}

int ref_rotate(RefVol const* dst, RefVol const* src, float const* axis, float angle, float const* center)
{
    if (dst->data == src->data) return -1;
    Staged d(dst), s(src);
    vkt::Rotate(d.vol, s.vol, { axis[0], axis[1], axis[2] }, angle, { center[0], center[1], center[2] });
    d.writeBack();
    return 0;
}

int ref_apply_filter(RefVol const* dst, RefVol const* src, float const* weights, int32_t const* fdims)
{
    for (int a = 0; a < 3; ++a)
        if (fdims[a] < 1 || fdims[a] > 3) return -1;       // the reference would read outside the volume
    std::vector<float> w(weights, weights + (size_t)fdims[0] * fdims[1] * fdims[2]);
    vkt::Filter filter(w.data(), v3(fdims));
    if (dst->data == src->data) {
        Staged v(dst);
        vkt::ApplyFilter(v.vol, v.vol, filter, vkt::AddressMode::Clamp);
        v.writeBack();
        return 0;
    }
    Staged d(dst), s(src);
    vkt::ApplyFilter(d.vol, s.vol, filter, vkt::AddressMode::Clamp);
    d.writeBack();
    return 0;
}

// vkt::OutputStream::writeRange / vkt::InputStream::readRange over a vkt::RawFile (include/cpp/vkt/OutputStream.hpp,
// InputStream.hpp, RawFile.hpp -> src/vkt/OutputStream.cpp:38-71, InputStream.cpp:41-72, RawFile.cpp:42-135).  The
// whole-volume read()/write() are the same calls with one line = the volume.  Returns the vkt::Error of the call.
int ref_write_range(RefVol const* vol, int32_t const* first, int32_t const* last, char const* path)
{
    Staged v(vol);
    vkt::RawFile file(path, "wb");
    vkt::OutputStream os(file);
    int e = (int)os.writeRange(v.vol, v3(first), v3(last));
    os.flush();
    return e;
}

int ref_read_range(RefVol const* vol, int32_t const* first, int32_t const* last, char const* path)
{
    Staged v(vol);
    vkt::RawFile file(path, "rb");
    vkt::InputStream is(file);
    int e = (int)is.readRange(v.vol, v3(first), v3(last));
    v.writeBack();
    return e;
}

int ref_write(RefVol const* vol, char const* path)
{
    Staged v(vol);
    vkt::RawFile file(path, "wb");
    vkt::OutputStream os(file);
    int e = (int)os.write(v.vol);
    os.flush();
    return e;
}

int ref_read(RefVol const* vol, char const* path)
{
    Staged v(vol);
    vkt::RawFile file(path, "rb");
    vkt::InputStream is(file);
    int e = (int)is.read(v.vol);
    v.writeBack();
    return e;
}

void ref_rawfile_parse(char const* name, int32_t* dims, int32_t* format)
{
    vkt::RawFile file(name, "rb");      // opens nothing when the file does not exist; the parser runs regardless
    vkt::Vec3i d = file.getDims();
    dims[0] = d.x; dims[1] = d.y; dims[2] = d.z;
    *format = (int32_t)file.getDataFormat();
}

// vkt::ComputeAggregatesRange (include/cpp/vkt/Aggregates.hpp:27-43 -> src/vkt/Aggregates.cpp:52-66).  out = the 52
// bytes of vkt::Aggregates.
int ref_aggregates_range(RefVol const* vol, int32_t const* first, int32_t const* last, void* out)
{
    for (int a = 0; a < 3; ++a)
        if (first[a] > last[a]) return -1;
    Staged v(vol);
    vkt::Aggregates ag;
    vkt::ComputeAggregatesRange(v.vol, ag, v3(first), v3(last));
    static_assert(sizeof(ag) == 52, "vkt::Aggregates layout");
    std::memcpy(out, &ag, sizeof(ag));
    return 0;
}

// vkt::ScanRange (include/cpp/vkt/Scan.hpp -> src/vkt/Scan.cpp:53-71).  Same data pointer = the in-place call on ONE
// volume object (the reference tests &dst != &src).
int ref_scan_range(RefVol const* dst, RefVol const* src, int32_t const* first, int32_t const* last, int32_t const* off)
{
    for (int a = 0; a < 3; ++a) {
        if (first[a] == last[a]) return -1;
        if (dst->data != src->data && first[a] > last[a]) return -1;
    }
    if (dst->data == src->data) {
        Staged v(dst);
        vkt::ScanRange(v.vol, v.vol, v3(first), v3(last), v3(off));
        v.writeBack();
        return 0;
    }
    Staged d(dst), s(src);
    vkt::ScanRange(d.vol, s.vol, v3(first), v3(last), v3(off));
    d.writeBack();
    return 0;
}

// vktBrickDecomposeResizeSV + vktBrickDecomposeSV (include/c/vkt/Decompose.h -> src/vkt/Decompose.cpp:152-256 ->
// BrickDecomposeC_serial, src/vkt/Decompose_serial.hpp:50-88).  The C entry points are used because the C++
// BrickDecomposeResize move-assigns volumes onto the malloc'ed, never-constructed elements of Array3D (its
// ManagedBuffer hands out raw bytes): whether that frees garbage depends on the heap's history.  The C array holds
// plain handles, so the same loops run without that hazard.
struct CSource
{
    vktStructuredVolume h;
    explicit CSource(RefVol const* d)
    {
        vktStructuredVolumeCreate(&h, d->dims[0], d->dims[1], d->dims[2], (vktDataFormat)d->format, 1.f, 1.f, 1.f,
                                  d->lo, d->hi);
        size_t bytes = h->volume.getSizeInBytes();   // the C accessors are declared but not defined by the reference
        if (bytes && d->data) std::memcpy(h->volume.getData(), d->data, bytes);
    }
    ~CSource() { vktStructuredVolumeDestroy(h); }
};

struct CBricks
{
    vktArray3D_vktStructuredVolume arr;
    CBricks(vktStructuredVolume source, int32_t const* b, int32_t const* n, int32_t const* p)
    {
        vktArray3D_vktStructuredVolume_CreateEmpty(&arr);
        vktBrickDecomposeResizeSV(arr, source, b[0], b[1], b[2], n[0], n[1], n[2], p[0], p[1], p[2]);
    }
    ~CBricks()
    {
        size_t n = vktArray3D_vktStructuredVolume_NumElements(arr);
        vktStructuredVolume* v = vktArray3D_vktStructuredVolume_Data(arr);
        for (size_t i = 0; i < n; ++i) vktStructuredVolumeDestroy(v[i]);
        vktArray3D_vktStructuredVolume_Destroy(arr);
    }
};

size_t ref_brick_layout(int32_t const* srcDims, int32_t const* brick, int32_t const* haloNeg, int32_t const* haloPos,
                        int32_t* numBricks, int32_t* brickDims)
{
    RefVol d = { nullptr, { srcDims[0], srcDims[1], srcDims[2] }, (int32_t)vkt::DataFormat::UInt8, 0.f, 1.f };
    CSource source(&d);
    CBricks bricks(source.h, brick, haloNeg, haloPos);
    vktVec3i_t nb = vktArray3D_vktStructuredVolume_Dims(bricks.arr);
    numBricks[0] = nb.x; numBricks[1] = nb.y; numBricks[2] = nb.z;
    size_t n = vktArray3D_vktStructuredVolume_NumElements(bricks.arr), total = 0;
    vktStructuredVolume* v = vktArray3D_vktStructuredVolume_Data(bricks.arr);
    for (size_t i = 0; i < n; ++i) {
        vkt::Vec3i bd = v[i]->volume.getDims();
        if (brickDims) { brickDims[3 * i] = bd.x; brickDims[3 * i + 1] = bd.y; brickDims[3 * i + 2] = bd.z; }
        total += (size_t)bd.x * (size_t)bd.y * (size_t)bd.z;
    }
    return total;
}

int ref_brick_decompose(RefVol const* src, int32_t const* brick, int32_t const* haloNeg, int32_t const* haloPos,
                        uint8_t* out, size_t outBytes)
{
    CSource source(src);
    CBricks bricks(source.h, brick, haloNeg, haloPos);
    vktBrickDecomposeSV(bricks.arr, source.h, brick[0], brick[1], brick[2], haloNeg[0], haloNeg[1], haloNeg[2],
                        haloPos[0], haloPos[1], haloPos[2]);
    size_t n = vktArray3D_vktStructuredVolume_NumElements(bricks.arr), total = 0;
    vktStructuredVolume* v = vktArray3D_vktStructuredVolume_Data(bricks.arr);
    for (size_t i = 0; i < n; ++i) total += v[i]->volume.getSizeInBytes();
    if (total != outBytes) return -2;
    for (size_t i = 0; i < n; ++i) {
        size_t bytes = v[i]->volume.getSizeInBytes();
        std::memcpy(out, v[i]->volume.getData(), bytes);
        out += bytes;
    }
    return 0;
}

} // extern "C"
