// volkit-b200 host mirror of the reference's public C++ API for the StructuredVolume hot path.
// Same names, argument meaning and error enum as reference include/cpp/vkt/common.hpp:14-66 (only the
// parts the hot path uses: Bool, Error, DataFormat).
#pragma once

#include <cstddef>
#include <cstdint>

#if defined(__GNUC__)
#define VKTAPI __attribute__((visibility("default")))
#else
#define VKTAPI
#endif

namespace vkt
{
    typedef uint8_t Bool;
    static constexpr Bool False = 0;
    static constexpr Bool True  = 1;

    // numeric values as in the reference (common.hpp:20-27)
    enum Error
    {
        InvalidValue      = -1,
        NoError           =  0,
        InvalidDataSource =  1,
        ReadError         =  2,
        WriteError        =  3,
    };

    // numeric values as in the reference (common.hpp:53-66); they are also the C-ABI's vktb_format values
    enum class DataFormat
    {
        Unspecified,
        Int8,
        Int16,
        Int32,
        UInt8,
        UInt16,
        UInt32,
        Float32,
        Count,   // keep last
    };

    class StructuredVolume;

    // Data source base class for file I/O (reference include/cpp/vkt/common.hpp:81-91)
    class DataSource
    {
    public:
        virtual ~DataSource() {}
        virtual std::size_t read(char* buf, std::size_t len) = 0;
        virtual std::size_t write(char const* buf, std::size_t len) = 0;
        virtual bool seek(std::size_t pos) = 0;
        virtual bool flush() = 0;
        virtual bool good() const = 0;
    };

} // vkt
