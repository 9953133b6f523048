// TEST INFRASTRUCTURE ONLY.  Link stubs for the two things the reference's command-line tool (src/cli/main.cpp,
// compiled where it lies) references but this box cannot provide:
//   * vkt::Render   -- the viewer needs a window system and OpenGL: unavailable on this box;
//   * vkt::VolumeFile -- src/vkt/VolumeFile.cpp needs boost::filesystem, which is not installed.
// Neither is on the path of BASELINE config 1 (declare-sv | fill | fill-range | resample over the stdin/stdout pipe), which is
// what oracle/_ref/vkt_ref_cli exists for: its bytes pin the pipe format our own `vkt` tool must write.
#include <cstddef>
#include <cstdio>

#include <vkt/InputStream.hpp>
#include <vkt/Render.hpp>
#include <vkt/VolumeFile.hpp>

namespace vkt
{
    Error Render(StructuredVolume&, RenderState const&, RenderState*)
    {
        std::fprintf(stderr, "vkt_ref_cli: render is unavailable on this box (no graphics stack)\n");
        return InvalidValue;
    }

    VolumeFile::VolumeFile(char const* fileName, OpenMode) : fileName_(fileName)
    {
        std::fprintf(stderr, "vkt_ref_cli: volume files are not built in (boost::filesystem is absent)\n");
    }
    VolumeFile::~VolumeFile() {}
    std::size_t VolumeFile::read(char*, std::size_t) { return 0; }
    std::size_t VolumeFile::write(char const*, std::size_t) { return 0; }
    bool VolumeFile::seek(std::size_t) { return false; }
    bool VolumeFile::flush() { return false; }
    bool VolumeFile::good() const { return false; }
    void VolumeFile::setHeader(VolumeFileHeader header) { header_ = header; }
    VolumeFileHeader VolumeFile::getHeader() const { return header_; }
} // namespace vkt
