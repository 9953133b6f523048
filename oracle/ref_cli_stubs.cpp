// TEST INFRASTRUCTURE ONLY.  Link stub for the one thing the reference's command-line tool (src/cli/main.cpp, compiled
// where it lies) references but this box cannot provide: vkt::Render -- the viewer needs a window system and OpenGL:
// unavailable on this box.  (src/vkt/VolumeFile.cpp is compiled unmodified against oracle/boost_stub/.)
// oracle/_ref/vkt_ref_cli exists to pin the bytes our own `vkt` tool must write.
#include <cstddef>
#include <cstdio>

#include <vkt/InputStream.hpp>
#include <vkt/Render.hpp>

namespace vkt
{
    Error Render(StructuredVolume&, RenderState const&, RenderState*)
    {
        std::fprintf(stderr, "vkt_ref_cli: render is unavailable on this box (no graphics stack)\n");
        return InvalidValue;
    }
} // namespace vkt
