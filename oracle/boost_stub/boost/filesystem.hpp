// TEST INFRASTRUCTURE ONLY.  A stand-in for <boost/filesystem.hpp>, which is not installed here: exactly what the
// reference's src/vkt/VolumeFile.cpp uses of it (path, extension(), change_extension, comparison with std::string), so
// that the file compiles unmodified into oracle/_ref/vkt_ref_cli and the tool's `read` command runs for real.
#pragma once

#include <string>

namespace boost
{
namespace filesystem
{
    class path
    {
    public:
        path(char const* s) : s_(s ? s : "") {}
        path(std::string s) : s_(std::move(s)) {}

        // ".raw" of "dir/name.raw"; empty when the file name has no dot (or only a leading one)
        path extension() const
        {
            std::size_t slash = s_.find_last_of('/');
            std::size_t start = slash == std::string::npos ? 0 : slash + 1;
            std::size_t dot = s_.find_last_of('.');
            if (dot == std::string::npos || dot <= start) return path(std::string());
            return path(s_.substr(dot));
        }

        std::string const& string() const { return s_; }

    private:
        std::string s_;
    };

    inline bool operator==(std::string const& a, path const& b) { return a == b.string(); }
    inline bool operator==(path const& a, std::string const& b) { return a.string() == b; }

    inline path change_extension(path const& p, std::string const& ext)
    {
        std::string s = p.string();
        std::size_t n = p.extension().string().size();
        return path(s.substr(0, s.size() - n) + ext);
    }
} // filesystem
} // boost
