// Raw volume files (reference include/cpp/vkt/RawFile.hpp:14-57, src/vkt/RawFile.cpp:42-175): a FILE* data source
// whose constructor parses "<dimx>x<dimy>x<dimz>" and "int<bits>" / "uint<bits>" tokens out of the '_'-separated
// file name.  read() / write() return fread's / fwrite's ITEM count for one item of `len` bytes (1 or 0), as the
// reference does -- InputStream::read therefore reports ReadError for every volume larger than one byte although the
// bytes arrive; callers of the reference ignore that code and so may callers of this mirror.
// RawFile(FILE*) adopts the stream (the reference's constructor fopen()s a null name there).
#pragma once

#include <cstddef>
#include <cstdio>

#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    class VKTAPI RawFile : public DataSource
    {
    public:
        RawFile(char const* fileName, char const* mode);
        RawFile(FILE* file);
       ~RawFile();

        virtual std::size_t read(char* buf, std::size_t len);
        virtual std::size_t write(char const* buf, std::size_t len);
        virtual bool seek(std::size_t pos);
        virtual bool flush();
        virtual bool good() const;

        void setDims(Vec3i dims);
        Vec3i getDims() const;
        void setDataFormat(DataFormat dataFormat);
        DataFormat getDataFormat() const;

    private:
        char const* fileName_ = 0;
        char const* mode_ = 0;
        FILE* file_ = 0;
        bool owned_ = true;

        Vec3i dims_ = { 0, 0, 0 };
        DataFormat dataFormat_ = DataFormat::UInt8;
    };
} // vkt
