// RawFile, InputStream, OutputStream (reference src/vkt/RawFile.cpp:42-175,
// src/vkt/InputStream.cpp:22-96, src/vkt/OutputStream.cpp:20-88).  CPU policy: the reference's statements.  GPU policy:
// chunked streaming through the backend's pinned staging ring (vktb_stream_in / vktb_stream_out).
#include <cstdio>
#include <cstring>
#include <sstream>
#include <string>
#include <vector>

#include <vkt/InputStream.hpp>
#include <vkt/OutputStream.hpp>
#include <vkt/RawFile.hpp>
#include <vkt/StructuredVolume.hpp>

#include "backend_link.hpp"
#include "host_codec.hpp"

namespace vkt
{
    //--- RawFile -------------------------------------------------------------------------------------------------
    RawFile::RawFile(char const* fileName, char const* mode) : fileName_(fileName), mode_(mode)
    {
        file_ = std::fopen(fileName_, mode_);
        // dimensions and voxel type from the '_'-separated tokens of the file name (RawFile.cpp:50-103)
        std::istringstream stream(fileName_);
        for (std::string token; std::getline(stream, token, '_'); ) {
            int32_t dimx = 0, dimy = 0, dimz = 0;
            uint16_t bpv = 0;
            if (std::sscanf(token.c_str(), "%dx%dx%d", &dimx, &dimy, &dimz) == 3) dims_ = { dimx, dimy, dimz };
            if (std::sscanf(token.c_str(), "int%hu", &bpv) == 1)
                dataFormat_ = bpv == 8 ? DataFormat::Int8 : bpv == 16 ? DataFormat::Int16
                            : bpv == 32 ? DataFormat::Int32 : DataFormat::Unspecified;
            if (std::sscanf(token.c_str(), "uint%hu", &bpv) == 1)
                dataFormat_ = bpv == 8 ? DataFormat::UInt8 : bpv == 16 ? DataFormat::UInt16
                            : bpv == 32 ? DataFormat::UInt32 : DataFormat::Unspecified;
        }
    }

    RawFile::RawFile(FILE* file) : file_(file), owned_(false) {}

    RawFile::~RawFile()
    {
        if (file_ && owned_) std::fclose(file_);
    }

    std::size_t RawFile::read(char* buf, std::size_t len) { return std::fread(buf, len, 1, file_); }
    std::size_t RawFile::write(char const* buf, std::size_t len) { return std::fwrite(buf, len, 1, file_); }
    bool RawFile::seek(std::size_t pos) { return good() && std::fseek(file_, (long)pos, SEEK_SET) == 0; }
    bool RawFile::flush() { return good() && std::fflush(file_) == 0; }
    bool RawFile::good() const { return file_ != nullptr; }
    void RawFile::setDims(Vec3i dims) { dims_ = dims; }
    Vec3i RawFile::getDims() const { return dims_; }
    void RawFile::setDataFormat(DataFormat dataFormat) { dataFormat_ = dataFormat; }
    DataFormat RawFile::getDataFormat() const { return dataFormat_; }

    namespace
    {
        size_t readTrampoline(void* ctx, void* buf, size_t len) { return ((DataSource*)ctx)->read((char*)buf, len); }
        size_t writeTrampoline(void* ctx, const void* buf, size_t len) { return ((DataSource*)ctx)->write((char const*)buf, len); }

        // the rows of [first, last) as the reference addresses them: lineBytes from the START of row (y, z)
        struct RowBox
        {
            int32_t rowVoxels, ny, nz;
            std::size_t lineBytes;
        };
        RowBox rowBox(StructuredVolume& v, int32_t fx, int32_t fy, int32_t fz, int32_t lx, int32_t ly, int32_t lz)
        {
            RowBox b;
            b.rowVoxels = lx - fx; b.ny = ly - fy; b.nz = lz - fz;
            b.lineBytes = (std::size_t)(lx - fx) * v.getBytesPerVoxel();
            return b;
        }
        vktb_vol boxVolume(void* data, const RowBox& b, StructuredVolume& like)
        {
            vktb_vol t = host::describe(like);
            t.data = data;
            t.dims[0] = b.rowVoxels; t.dims[1] = b.ny; t.dims[2] = b.nz;
            return t;
        }
    }

    //--- InputStream ---------------------------------------------------------------------------------------------
    InputStream::InputStream(DataSource& source) : dataSource_(source) {}

    Error InputStream::read(StructuredVolume& volume)
    {
        if (!dataSource_.good()) return InvalidDataSource;
        const std::size_t bytes = volume.getSizeInBytes();
        std::size_t len = 0;
        if (host::onGpu(GetThreadExecutionPolicy())) {
            volume.migrateDiscard();
            if (vktb_stream_in(volume.dataInPlace(), bytes, readTrampoline, &dataSource_, &len) != 0)
                return host::backendError("InputStream::read");
        } else {
            len = dataSource_.read((char*)volume.getData(), bytes);
        }
        return len != bytes ? ReadError : NoError;
    }

    Error InputStream::readRange(StructuredVolume& dst, int32_t firstX, int32_t firstY, int32_t firstZ,
                                 int32_t lastX, int32_t lastY, int32_t lastZ)
    {
        const uint8_t bpv = dst.getBytesPerVoxel();
        const Vec3i dims = dst.getDims();
        const RowBox b = rowBox(dst, firstX, firstY, firstZ, lastX, lastY, lastZ);
        if (b.rowVoxels < 0 || b.ny < 0 || b.nz < 0) return InvalidValue;           // the reference's != loops never end
        const std::size_t want = (std::size_t)b.nz * b.ny * b.lineBytes;
        std::size_t len = 0;
        if (host::onGpu(GetThreadExecutionPolicy())) {
            // the rows arrive back to back in a device scratch box and one CopyRange moves them to the row starts
            if (want == 0) return NoError;
            void* box = nullptr;
            if (vktb_alloc(&box, want) != 0) return host::backendError("InputStream::readRange");
            struct Lines { DataSource* src; std::size_t line; } ctx{ &dataSource_, b.lineBytes };
            auto readLines = [](void* c, void* buf, size_t n) -> size_t {
                Lines* l = (Lines*)c;
                size_t got = 0;
                for (size_t o = 0; o < n; ) {               // whole rows, except where a chunk border cuts one
                    const size_t part = std::min(l->line, n - o);
                    got += l->src->read((char*)buf + o, part);
                    o += part;
                }
                return got;
            };
            // chunk borders (16 MiB) may cut a row in two reads; sources that count items then report one item more
            int rc = vktb_stream_in(box, want, readLines, &ctx, &len);
            vktb_vol d = host::describe(dst), t = boxVolume(box, b, dst);
            const int32_t z3[3] = { 0, 0, 0 }, l3[3] = { b.rowVoxels, b.ny, b.nz }, off[3] = { 0, firstY, firstZ };
            if (rc == 0) rc = vktb_copy_range(&d, &t, z3, l3, off);
            vktb_free(box);
            if (rc != 0) return host::backendError("InputStream::readRange");
        } else {
            for (int32_t z = firstZ; z != lastZ; ++z)
                for (int32_t y = firstY; y != lastY; ++y) {
                    const std::size_t offset = (z * dims.x * std::size_t(dims.y) + y * dims.x) * bpv;
                    len += dataSource_.read((char*)dst.getData() + offset, b.lineBytes);
                }
        }
        return len != want ? ReadError : NoError;
    }

    Error InputStream::readRange(StructuredVolume& dst, Vec3i first, Vec3i last)
    {
        return readRange(dst, first.x, first.y, first.z, last.x, last.y, last.z);
    }

    Error InputStream::seek(std::size_t pos) { return dataSource_.seek(pos) ? NoError : InvalidValue; }

    //--- OutputStream --------------------------------------------------------------------------------------------
    OutputStream::OutputStream(DataSource& source) : dataSource_(source) {}

    Error OutputStream::write(StructuredVolume& volume)
    {
        if (!dataSource_.good()) return InvalidDataSource;
        const std::size_t bytes = volume.getSizeInBytes();
        std::size_t len = 0;
        if (host::onGpu(GetThreadExecutionPolicy())) {
            if (vktb_stream_out(volume.getData(), bytes, writeTrampoline, &dataSource_, &len) != 0)
                return host::backendError("OutputStream::write");
        } else {
            len = dataSource_.write((char const*)volume.getData(), bytes);
        }
        return len != bytes ? WriteError : NoError;
    }

    Error OutputStream::writeRange(StructuredVolume& dst, int32_t firstX, int32_t firstY, int32_t firstZ,
                                   int32_t lastX, int32_t lastY, int32_t lastZ)
    {
        const uint8_t bpv = dst.getBytesPerVoxel();
        const Vec3i dims = dst.getDims();
        const RowBox b = rowBox(dst, firstX, firstY, firstZ, lastX, lastY, lastZ);
        if (b.rowVoxels < 0 || b.ny < 0 || b.nz < 0) return InvalidValue;
        const std::size_t want = (std::size_t)b.nz * b.ny * b.lineBytes;
        std::size_t len = 0;
        if (host::onGpu(GetThreadExecutionPolicy())) {
            if (want == 0) return NoError;
            void* box = nullptr;
            if (vktb_alloc(&box, want) != 0) return host::backendError("OutputStream::writeRange");
            vktb_vol s = host::describe(dst), t = boxVolume(box, b, dst);
            const int32_t f3[3] = { 0, firstY, firstZ }, l3[3] = { b.rowVoxels, lastY, lastZ }, off[3] = { 0, 0, 0 };
            int rc = vktb_copy_range(&t, &s, f3, l3, off);
            struct Lines { DataSource* src; std::size_t line; } ctx{ &dataSource_, b.lineBytes };
            auto writeLines = [](void* c, const void* buf, size_t n) -> size_t {
                Lines* l = (Lines*)c;
                size_t put = 0;
                for (size_t o = 0; o < n; ) {
                    const size_t part = std::min(l->line, n - o);
                    put += l->src->write((char const*)buf + o, part);
                    o += part;
                }
                return put;
            };
            if (rc == 0) rc = vktb_stream_out(box, want, writeLines, &ctx, &len);
            vktb_free(box);
            if (rc != 0) return host::backendError("OutputStream::writeRange");
        } else {
            for (int32_t z = firstZ; z != lastZ; ++z)
                for (int32_t y = firstY; y != lastY; ++y) {
                    const std::size_t offset = (z * dims.x * std::size_t(dims.y) + y * dims.x) * bpv;
                    len += dataSource_.write((char const*)dst.getData() + offset, b.lineBytes);
                }
        }
        return len != want ? WriteError : NoError;
    }

    Error OutputStream::writeRange(StructuredVolume& dst, Vec3i first, Vec3i last)
    {
        return writeRange(dst, first.x, first.y, first.z, last.x, last.y, last.z);
    }

    Error OutputStream::seek(std::size_t pos) { return dataSource_.seek(pos) ? NoError : InvalidValue; }
    Error OutputStream::flush() { return dataSource_.flush() ? NoError : InvalidValue; }
} // vkt
