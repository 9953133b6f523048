// StructuredVolume (reference src/vkt/StructuredVolume.cpp:24-322) and the scalar codec entry points
// (reference src/vkt/Voxel.cpp).
#include <cstring>

#include <vkt/StructuredVolume.hpp>
#include <vkt/Voxel.hpp>

#include "backend_link.hpp"
#include "host_codec.hpp"

namespace vkt
{
    static std::size_t volumeBytes(Vec3i d, DataFormat f)
    {
        return (std::size_t)d.x * (std::size_t)d.y * (std::size_t)d.z * (std::size_t)host::bytesPerVoxel(f);
    }

    StructuredVolume::StructuredVolume()
        : ManagedBuffer(0), dims_{ 0, 0, 0 }, dataFormat_(DataFormat::UInt8), dist_{ 1.f, 1.f, 1.f },
          voxelMapping_{ 0.f, 1.f }, haloSize_{ .5f, .5f, .5f }
    {
    }

    StructuredVolume::StructuredVolume(int32_t dimX, int32_t dimY, int32_t dimZ, DataFormat dataFormat, float distX,
                                       float distY, float distZ, float mappingLo, float mappingHi)
        : ManagedBuffer(volumeBytes({ dimX, dimY, dimZ }, dataFormat)), dims_{ dimX, dimY, dimZ },
          dataFormat_(dataFormat), dist_{ distX, distY, distZ }, voxelMapping_{ mappingLo, mappingHi },
          haloSize_{ .5f, .5f, .5f }
    {
    }

    void StructuredVolume::setDims(int32_t dimX, int32_t dimY, int32_t dimZ) { setDims(Vec3i{ dimX, dimY, dimZ }); }
    void StructuredVolume::getDims(int32_t& dimX, int32_t& dimY, int32_t& dimZ) { dimX = dims_.x; dimY = dims_.y; dimZ = dims_.z; }
    void StructuredVolume::setDims(Vec3i dims) { dims_ = dims; resize(getSizeInBytes()); }
    Vec3i StructuredVolume::getDims() const { return dims_; }

    void StructuredVolume::setDataFormat(DataFormat dataFormat) { dataFormat_ = dataFormat; resize(getSizeInBytes()); }
    DataFormat StructuredVolume::getDataFormat() const { return dataFormat_; }

    void StructuredVolume::setDist(float distX, float distY, float distZ) { dist_ = { distX, distY, distZ }; }
    void StructuredVolume::getDist(float& distX, float& distY, float& distZ) { distX = dist_.x; distY = dist_.y; distZ = dist_.z; }
    void StructuredVolume::setDist(Vec3f dist) { dist_ = dist; }
    Vec3f StructuredVolume::getDist() const { return dist_; }

    void StructuredVolume::setVoxelMapping(float lo, float hi) { voxelMapping_ = { lo, hi }; }
    void StructuredVolume::getVoxelMapping(float& lo, float& hi) { lo = voxelMapping_.x; hi = voxelMapping_.y; }
    void StructuredVolume::setVoxelMapping(Vec2f mapping) { voxelMapping_ = mapping; }
    Vec2f StructuredVolume::getVoxelMapping() const { return voxelMapping_; }

    Box3f StructuredVolume::getObjectBounds() const
    {
        return { { 0.f, 0.f, 0.f }, { dims_.x * dist_.x, dims_.y * dist_.y, dims_.z * dist_.z } };
    }

    Box3f StructuredVolume::getDomainBounds() const
    {
        Box3f b = getObjectBounds();
        b.min = { b.min.x - haloSize_.x, b.min.y - haloSize_.y, b.min.z - haloSize_.z };
        b.max = { b.max.x + haloSize_.x, b.max.y + haloSize_.y, b.max.z + haloSize_.z };
        return b;
    }

    uint8_t* StructuredVolume::getData()
    {
        migrate();
        return data_;
    }

    // Accessors migrate first (reference :160-296).  After that the bytes are where the thread's policy says: host
    // memory is dereferenced directly; under Device::GPU one voxel travels through vktb_memcpy (the reference
    // dereferences the device pointer on the host at this point, which cannot work).
    void StructuredVolume::setBytes(int32_t x, int32_t y, int32_t z, uint8_t const* data)
    {
        migrate();
        const std::size_t off = linearIndex(x, y, z);
        const std::size_t n = getBytesPerVoxel();
        if (host::onGpu(GetThreadExecutionPolicy())) {
            if (vktb_memcpy(data_ + off, data, n, VKTB_H2D) != 0) host::backendError("StructuredVolume::setBytes");
        } else {
            std::memcpy(data_ + off, data, n);
        }
    }

    void StructuredVolume::getBytes(int32_t x, int32_t y, int32_t z, uint8_t* data)
    {
        migrate();
        const std::size_t off = linearIndex(x, y, z);
        const std::size_t n = getBytesPerVoxel();
        if (host::onGpu(GetThreadExecutionPolicy())) {
            if (vktb_memcpy(data, data_ + off, n, VKTB_D2H) != 0) host::backendError("StructuredVolume::getBytes");
        } else {
            std::memcpy(data, data_ + off, n);
        }
    }

    void StructuredVolume::setValue(int32_t x, int32_t y, int32_t z, float value)
    {
        uint8_t raw[8] = {};
        host::mapVoxel(raw, value, dataFormat_, voxelMapping_.x, voxelMapping_.y);
        setBytes(x, y, z, raw);
    }

    void StructuredVolume::getValue(int32_t x, int32_t y, int32_t z, float& value)
    {
        uint8_t raw[8] = {};
        getBytes(x, y, z, raw);
        value = host::unmapVoxel(raw, dataFormat_, voxelMapping_.x, voxelMapping_.y);
    }

    float StructuredVolume::getValue(int32_t x, int32_t y, int32_t z)
    {
        float v = 0.f;
        getValue(x, y, z, v);
        return v;
    }

    void StructuredVolume::setValue(Vec3i i, float value) { setValue(i.x, i.y, i.z, value); }
    void StructuredVolume::getValue(Vec3i i, float& value) { getValue(i.x, i.y, i.z, value); }
    float StructuredVolume::getValue(Vec3i i) { return getValue(i.x, i.y, i.z); }
    void StructuredVolume::setBytes(Vec3i i, uint8_t const* data) { setBytes(i.x, i.y, i.z, data); }
    void StructuredVolume::getBytes(Vec3i i, uint8_t* data) { getBytes(i.x, i.y, i.z, data); }

    uint8_t StructuredVolume::getBytesPerVoxel() const { return (uint8_t)host::bytesPerVoxel(dataFormat_); }
    std::size_t StructuredVolume::getSizeInBytes() const { return volumeBytes(dims_, dataFormat_); }

    std::size_t StructuredVolume::linearIndex(int32_t x, int32_t y, int32_t z) const
    {
        return ((std::size_t)z * dims_.x * (std::size_t)dims_.y + (std::size_t)y * dims_.x + (std::size_t)x) * getBytesPerVoxel();
    }

    Error MapVoxel(uint8_t* dst, float value, DataFormat dataFormat, float mappingLo, float mappingHi)
    {
        host::mapVoxel(dst, value, dataFormat, mappingLo, mappingHi);
        return NoError;
    }

    Error UnmapVoxel(float& value, uint8_t const* src, DataFormat dataFormat, float mappingLo, float mappingHi)
    {
        value = host::unmapVoxel(src, dataFormat, mappingLo, mappingHi);
        return NoError;
    }
} // vkt
