// Managed structured volume (reference include/cpp/vkt/StructuredVolume.hpp:34-132,
// src/vkt/StructuredVolume.cpp:24-322): dense x-fastest grid, byte offset ((z*dimY + y)*dimX + x) * bytesPerVoxel,
// DataFormat voxels with a linear [lo,hi] voxel mapping.  Where the bytes live (host or B200) follows the
// calling thread's ExecutionPolicy, lazily: every accessor and algorithm migrates first.
#pragma once

#include <cstddef>
#include <cstdint>

#include "ManagedBuffer.hpp"
#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    class VKTAPI StructuredVolume : public ManagedBuffer<uint8_t>
    {
    public:
        constexpr static uint8_t GetMaxBytesPerVoxel() { return 8; }

        StructuredVolume();
        StructuredVolume(int32_t dimX, int32_t dimY, int32_t dimZ, DataFormat dataFormat,
                         float distX = 1.f, float distY = 1.f, float distZ = 1.f,
                         float mappingLo = 0.f, float mappingHi = 1.f);
        StructuredVolume(StructuredVolume& rhs) = default;
        StructuredVolume(StructuredVolume&& rhs) = default;
        StructuredVolume& operator=(StructuredVolume& rhs) = default;
        StructuredVolume& operator=(StructuredVolume&& rhs) = default;

        void setDims(int32_t dimX, int32_t dimY, int32_t dimZ);
        void getDims(int32_t& dimX, int32_t& dimY, int32_t& dimZ);
        void setDims(Vec3i dims);
        Vec3i getDims() const;

        void setDataFormat(DataFormat dataFormat);
        DataFormat getDataFormat() const;

        void setDist(float distX, float distY, float distZ);
        void getDist(float& distX, float& distY, float& distZ);
        void setDist(Vec3f dist);
        Vec3f getDist() const;

        void setVoxelMapping(float lo, float hi);
        void getVoxelMapping(float& lo, float& hi);
        void setVoxelMapping(Vec2f mapping);
        Vec2f getVoxelMapping() const;

        Box3f getDomainBounds() const;   // object bounds grown by the (fixed) half-voxel halo
        Box3f getObjectBounds() const;   // origin at (0,0,0), dims * dist

        // raw pointer to the data ON THE CURRENT POLICY'S DEVICE (migrates first)
        uint8_t* getData();

        void setValue(int32_t x, int32_t y, int32_t z, float value);
        void getValue(int32_t x, int32_t y, int32_t z, float& value);
        float getValue(int32_t x, int32_t y, int32_t z);
        void setValue(Vec3i index, float value);
        void getValue(Vec3i index, float& value);
        float getValue(Vec3i index);

        void setBytes(int32_t x, int32_t y, int32_t z, uint8_t const* data);
        void getBytes(int32_t x, int32_t y, int32_t z, uint8_t* data);
        void setBytes(Vec3i index, uint8_t const* data);
        void getBytes(Vec3i index, uint8_t* data);

        uint8_t getBytesPerVoxel() const;
        std::size_t getSizeInBytes() const;

    private:
        Vec3i dims_;
        DataFormat dataFormat_;
        Vec3f dist_;
        Vec2f voxelMapping_;
        Vec3f haloSize_;

        std::size_t linearIndex(int32_t x, int32_t y, int32_t z) const;
    };
} // vkt
