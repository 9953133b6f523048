// `vkt` command-line tool, hot-path subset: declare-sv | fill | fill-range | flip | flip-range | resample | dump | dump-range
// (+ set-header, an empty stub in the reference too), piping a structured
// volume over stdin/stdout in the reference's binary pipe format (reference src/cli/main.cpp:32-88):
//     u32 magic = 1, u32 assetType = 0, i32 dims[3], u32 dataFormat, f32 dist[3], f32 mapping[2], raw voxels
// Commands, options, defaults and range checks follow reference src/cli/main.cpp:111-335 (options), :402-442
// (declare-sv), :521-574 (fill / fill-range), :575-620 (flip / flip-range), :701-740 (resample, always
// FilterMode::Nearest), :443-519 (dump, also from a file with -i), :630-656 (read), :741-747 (set-header), :748-781 (write).
// Files: raw volumes only ("<name>_<X>x<Y>x<Z>_<format>.raw", vkt::RawFile); the reference's VolumeFile also knows Nifti and
// Virvo when built with those libraries.  Where the reference tool crashes -- a raw file that does not exist, and every `write`
// (its VolumeFile opens raw files read-only, src/vkt/VolumeFile.cpp:73) -- this tool reports the error / writes the file.
// render needs the reference's OpenGL viewer: not built.
//
// volkit-b200 addition: the execution policy of every stage is chosen with `--device cpu|gpu` or the environment
// variable VKT_B200_DEVICE (default cpu, as in the reference where the CLI never sets a policy).  Under the GPU
// policy the voxels of the pipe go through vkt::InputStream / vkt::OutputStream: stdin is read
// in 16 MiB pieces straight into the backend's pinned staging ring while the previous pieces are on their way to
// the device, and the result leaves the same way -- the stage never holds a host copy of the volume.
#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <vkt/ExecutionPolicy.hpp>
#include <vkt/Fill.hpp>
#include <vkt/Flip.hpp>
#include <vkt/InputStream.hpp>
#include <vkt/OutputStream.hpp>
#include <vkt/RawFile.hpp>
#include <vkt/Resample.hpp>
#include <vkt/StructuredVolume.hpp>

using namespace vkt;

namespace
{
struct Header
{
    uint32_t magic, assetType;
    int32_t  dims[3];
    uint32_t dataFormat;
    float    dist[3];
    float    mapping[2];
};
static_assert(sizeof(Header) == 44, "pipe header is 44 bytes");

struct Options
{
    std::string command;
    Vec3i dims{ 0, 0, 0 };
    DataFormat dataFormat = DataFormat::Unspecified;
    Vec3f dist{ 0.f, 0.f, 0.f };
    Vec2f mapping{ 0.f, 0.f };
    Vec3i first{ 0, 0, 0 }, last{ 0, 0, 0 };
    float value = 0.f;
    Axis axis = Axis::X;
    std::string inputFile, outputFile;
    std::string device;
};

bool fail(const char* msg) { std::fprintf(stderr, "%s\n", msg); return false; }

DataFormat parseFormat(std::string s)
{
    std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char)std::tolower(c); });
    static const struct { const char* name; DataFormat f; } table[] = {
        { "int8", DataFormat::Int8 }, { "int16", DataFormat::Int16 }, { "int32", DataFormat::Int32 },
        { "uint8", DataFormat::UInt8 }, { "uint16", DataFormat::UInt16 }, { "uint32", DataFormat::UInt32 },
        { "float32", DataFormat::Float32 } };
    for (auto& e : table) if (s == e.name) return e.f;
    return DataFormat::Unspecified;
}

bool parse(int argc, char** argv, Options& o)
{
    if (argc < 2) return fail("Usage: vkt <command> [<args>]");
    o.command = argv[1];
    static const char* known[] = { "declare-sv", "dump", "dump-range", "fill", "fill-range", "flip", "flip-range", "read", "resample",
                                   "set-header", "write" };
    if (std::find_if(std::begin(known), std::end(known), [&](const char* k) { return o.command == k; }) == std::end(known)) {
        std::fprintf(stderr, "Usage: vkt <command> [<args>]\n  Command \"%s\" unknown\n", argv[1]);
        return false;
    }
    auto need = [&](int i, int n, const char* opt) {
        if (i + n >= argc) { std::fprintf(stderr, "Option %s requires %d argument(s)\n", opt, n); return false; }
        return true;
    };
    for (int i = 2; i < argc; ++i) {
        std::string opt = argv[i];
        if (opt == "-df" || opt == "--data-format") { if (!need(i, 1, argv[i])) return false; o.dataFormat = parseFormat(argv[++i]); }
        else if (opt == "-dims" || opt == "--dims") { if (!need(i, 3, argv[i])) return false; o.dims = { std::atoi(argv[i + 1]), std::atoi(argv[i + 2]), std::atoi(argv[i + 3]) }; i += 3; }
        else if (opt == "-dist" || opt == "--dist") { if (!need(i, 3, argv[i])) return false; o.dist = { (float)std::atof(argv[i + 1]), (float)std::atof(argv[i + 2]), (float)std::atof(argv[i + 3]) }; i += 3; }
        else if (opt == "-first" || opt == "--first") { if (!need(i, 3, argv[i])) return false; o.first = { std::atoi(argv[i + 1]), std::atoi(argv[i + 2]), std::atoi(argv[i + 3]) }; i += 3; }
        else if (opt == "-last" || opt == "--last") { if (!need(i, 3, argv[i])) return false; o.last = { std::atoi(argv[i + 1]), std::atoi(argv[i + 2]), std::atoi(argv[i + 3]) }; i += 3; }
        else if (opt == "-vm" || opt == "--voxel-mapping") { if (!need(i, 2, argv[i])) return false; o.mapping = { (float)std::atof(argv[i + 1]), (float)std::atof(argv[i + 2]) }; i += 2; }
        else if (opt == "-val" || opt == "-v" || opt == "--value") { if (!need(i, 1, argv[i])) return false; o.value = (float)std::atof(argv[++i]); }
        else if (opt == "-ax" || opt == "--axis") {
            if (!need(i, 1, argv[i])) return false;
            const char* a = argv[++i];
            const int c = std::strlen(a) == 1 ? std::tolower((unsigned char)a[0]) : 0;
            if (c == 'x') o.axis = Axis::X; else if (c == 'y') o.axis = Axis::Y; else if (c == 'z') o.axis = Axis::Z;
            else { std::fprintf(stderr, "Invalid argument: %s\n", a); return false; }
        }
        else if (opt == "-i" || opt == "--input") { if (!need(i, 1, argv[i])) return false; o.inputFile = argv[++i]; }
        else if (opt == "-o" || opt == "--output") { if (!need(i, 1, argv[i])) return false; o.outputFile = argv[++i]; }
        else if (opt == "--device") { if (!need(i, 1, argv[i])) return false; o.device = argv[++i]; }
        else { std::fprintf(stderr, "Unknown option %s\n", argv[i]); return false; }
    }
    return true;
}

bool readAll(void* dst, size_t n) { return std::fread(dst, 1, n, stdin) == n; }

// stdin / stdout as a DataSource that counts bytes
class PipeSource : public DataSource
{
public:
    explicit PipeSource(FILE* f) : file_(f) {}
    std::size_t read(char* buf, std::size_t len) override { return std::fread(buf, 1, len, file_); }
    std::size_t write(char const* buf, std::size_t len) override { return std::fwrite(buf, 1, len, file_); }
    bool seek(std::size_t) override { return false; }
    bool flush() override { return std::fflush(file_) == 0; }
    bool good() const override { return file_ != nullptr; }
private:
    FILE* file_;
};

// reads header + voxels from stdin into a volume allocated under the calling thread's policy
bool readVolume(StructuredVolume& volume)
{
    Header h;
    if (!readAll(&h, sizeof h)) return fail("Cannot read structured volume: short header");
    if (h.magic != 1u) return fail("Cannot read structured volume, wrong version");
    if (h.assetType != 0u) return fail("Cannot read structured volume, wrong type");
    volume = StructuredVolume(h.dims[0], h.dims[1], h.dims[2], (DataFormat)h.dataFormat, h.dist[0], h.dist[1], h.dist[2],
                              h.mapping[0], h.mapping[1]);
    PipeSource in(stdin);
    if (InputStream(in).read(volume) != NoError) return fail("Cannot read structured volume: short data");
    return true;
}

// CPU policy: getData() migrates the result back to the host before it is written; GPU policy: streamed out
void writeVolume(StructuredVolume& volume)
{
    Header h;
    h.magic = 1u; h.assetType = 0u;
    Vec3i d = volume.getDims(); Vec3f s = volume.getDist(); Vec2f m = volume.getVoxelMapping();
    h.dims[0] = d.x; h.dims[1] = d.y; h.dims[2] = d.z;
    h.dataFormat = (uint32_t)volume.getDataFormat();
    h.dist[0] = s.x; h.dist[1] = s.y; h.dist[2] = s.z;
    h.mapping[0] = m.x; h.mapping[1] = m.y;
    std::fwrite(&h, 1, sizeof h, stdout);
    PipeSource out(stdout);
    OutputStream os(out);
    os.write(volume);
    os.flush();
}

// "<name>.raw" / ".RAW" (reference src/vkt/VolumeFile.cpp:24-33); the other formats need libraries that are not built in
bool isRawName(const std::string& name)
{
    const std::size_t slash = name.find_last_of('/'), dot = name.find_last_of('.');
    if (dot == std::string::npos || (slash != std::string::npos && dot < slash)) return false;
    const std::string ext = name.substr(dot);
    return ext == ".raw" || ext == ".RAW";
}

// a raw file into a volume allocated under the calling thread's policy (dims and format come from the file name)
bool readRawFile(const std::string& name, StructuredVolume& volume)
{
    FILE* probe = isRawName(name) ? std::fopen(name.c_str(), "rb") : nullptr;
    if (!probe) { std::fprintf(stderr, "Cannot open input file: %s\n", name.c_str()); return false; }
    std::fclose(probe);
    RawFile file(name.c_str(), "rb");
    Vec3i d = file.getDims();
    if ((long long)d.x * d.y * d.z < 1) { std::fprintf(stderr, "Invalid dimensions: %s\n", name.c_str()); return false; }
    if (file.getDataFormat() == DataFormat::Unspecified) { std::fprintf(stderr, "Invalid data format: %s\n", name.c_str()); return false; }
    volume = StructuredVolume(d.x, d.y, d.z, file.getDataFormat(), 1.f, 1.f, 1.f, 0.f, 1.f);
    InputStream(file).read(volume);     // RawFile::read returns an item count (RawFile.hpp): the code is not a byte count check
    return true;
}

bool checkParams(StructuredVolume& v)
{
    Vec3i d = v.getDims(); Vec3f s = v.getDist(); Vec2f m = v.getVoxelMapping();
    if ((long long)d.x * d.y * d.z < 1) return fail("Invalid dimensions");
    if (v.getDataFormat() == DataFormat::Unspecified) return fail("Invalid data format");
    if (s.x * s.y * s.z <= 0.f) return fail("Invalid dist");
    if (m.y - m.x <= 0.f) return fail("Invalid voxel mapping");
    return true;
}

ExecutionPolicy policyFor(const Options& o, bool gpu)
{
    ExecutionPolicy ep;
    (void)o;
    ep.device = gpu ? ExecutionPolicy::Device::GPU : ExecutionPolicy::Device::CPU;
    return ep;
}
} // namespace

int main(int argc, char** argv)
{
    Options o;
    if (!parse(argc, argv, o)) return EXIT_FAILURE;
    if (o.device.empty()) { const char* e = std::getenv("VKT_B200_DEVICE"); o.device = e ? e : "cpu"; }
    const bool gpu = o.device == "gpu" || o.device == "GPU";
    const ExecutionPolicy cpuPolicy = policyFor(o, false), runPolicy = policyFor(o, gpu);

    if (o.command == "declare-sv") {
        if ((long long)o.dims.x * o.dims.y * o.dims.z <= 0) { fail("Dims required"); return EXIT_FAILURE; }
        if (o.dataFormat == DataFormat::Unspecified) { fail("Data format required"); return EXIT_FAILURE; }
        Vec3f dist = o.dist.x * o.dist.y * o.dist.z <= 0.f ? Vec3f{ 1.f, 1.f, 1.f } : o.dist;
        Vec2f mapping = o.mapping.y - o.mapping.x <= 0.f ? Vec2f{ 0.f, 1.f } : o.mapping;
        StructuredVolume volume(o.dims.x, o.dims.y, o.dims.z, o.dataFormat, dist.x, dist.y, dist.z, mapping.x, mapping.y);
        // the reference emits uninitialised memory here; zeros are as good and reproducible
        std::memset(volume.getData(), 0, volume.getSizeInBytes());
        writeVolume(volume);
    } else if (o.command == "fill" || o.command == "fill-range") {
        SetThreadExecutionPolicy(runPolicy);          // the pipe is streamed to / from the device that computes
        StructuredVolume volume;
        if (!readVolume(volume) || !checkParams(volume)) return EXIT_FAILURE;
        const long long range = (long long)(o.last.x - o.first.x) * (o.last.y - o.first.y) * (o.last.z - o.first.z);
        if (o.command == "fill-range") {
            if (range <= 0) { fail("Invalid range"); return EXIT_FAILURE; }
            if (o.first.x < 0 || o.first.y < 0 || o.first.z < 0) { fail("Range underflow"); return EXIT_FAILURE; }
            Vec3i d = volume.getDims();
            if (o.last.x > d.x || o.last.y > d.y || o.last.z > d.z) { fail("Range overflow"); return EXIT_FAILURE; }
        }
        Error e = range > 0 ? FillRange(volume, o.first, o.last, o.value) : Fill(volume, o.value);
        if (e != NoError) return EXIT_FAILURE;
        writeVolume(volume);
        SetThreadExecutionPolicy(cpuPolicy);
    } else if (o.command == "flip" || o.command == "flip-range") {
        SetThreadExecutionPolicy(runPolicy);
        StructuredVolume source;
        if (!readVolume(source) || !checkParams(source)) return EXIT_FAILURE;
        StructuredVolume dest(source);
        const long long range = (long long)(o.last.x - o.first.x) * (o.last.y - o.first.y) * (o.last.z - o.first.z);
        if (o.command == "flip-range") {
            if (range <= 0) { fail("Invalid range"); return EXIT_FAILURE; }
            if (o.first.x < 0 || o.first.y < 0 || o.first.z < 0) { fail("Range underflow"); return EXIT_FAILURE; }
            Vec3i d = source.getDims();
            if (o.last.x > d.x || o.last.y > d.y || o.last.z > d.z) { fail("Range overflow"); return EXIT_FAILURE; }
        }
        Error e = range > 0 ? FlipRange(dest, source, o.first, o.last, o.axis) : Flip(dest, source, o.axis);
        if (e != NoError) return EXIT_FAILURE;
        writeVolume(dest);
        SetThreadExecutionPolicy(cpuPolicy);
    } else if (o.command == "read") {
        if (o.inputFile.empty()) { fail("Input file missing"); return EXIT_FAILURE; }
        SetThreadExecutionPolicy(runPolicy);
        StructuredVolume volume;
        if (!readRawFile(o.inputFile, volume)) return EXIT_FAILURE;
        writeVolume(volume);
        SetThreadExecutionPolicy(cpuPolicy);
    } else if (o.command == "write") {
        if (o.outputFile.empty()) { fail("Output file missing"); return EXIT_FAILURE; }
        StructuredVolume volume;
        if (!readVolume(volume) || !checkParams(volume)) return EXIT_FAILURE;
        if (!isRawName(o.outputFile)) { std::fprintf(stderr, "Cannot open output file: %s\n", o.outputFile.c_str()); return EXIT_FAILURE; }
        RawFile file(o.outputFile.c_str(), "wb");
        if (!file.good()) { std::fprintf(stderr, "Cannot open output file: %s\n", o.outputFile.c_str()); return EXIT_FAILURE; }
        OutputStream os(file);
        os.write(volume);
        os.flush();
    } else if (o.command == "set-header") {
        // the reference's command reads nothing and writes nothing (src/cli/main.cpp:741-747)
    } else if (o.command == "resample") {
        SetThreadExecutionPolicy(runPolicy);
        StructuredVolume src;
        if (!readVolume(src)) return EXIT_FAILURE;
        Vec3i dims = (long long)o.dims.x * o.dims.y * o.dims.z <= 0 ? src.getDims() : o.dims;
        DataFormat f = o.dataFormat == DataFormat::Unspecified ? src.getDataFormat() : o.dataFormat;
        Vec3f dist = o.dist.x * o.dist.y * o.dist.z <= 0.f ? src.getDist() : o.dist;
        Vec2f mapping = o.mapping.y - o.mapping.x <= 0.f ? src.getVoxelMapping() : o.mapping;
        StructuredVolume dst(dims.x, dims.y, dims.z, f, dist.x, dist.y, dist.z, mapping.x, mapping.y);
        Error e = Resample(dst, src, FilterMode::Nearest);
        if (e != NoError) return EXIT_FAILURE;
        writeVolume(dst);
        SetThreadExecutionPolicy(cpuPolicy);
    } else {   // dump / dump-range
        StructuredVolume volume;
        if (o.inputFile.empty() ? !readVolume(volume) : !readRawFile(o.inputFile, volume)) return EXIT_FAILURE;
        const long long range = (long long)(o.last.x - o.first.x) * (o.last.y - o.first.y) * (o.last.z - o.first.z);
        if (o.command == "dump-range" && range <= 0) { fail("Invalid range"); return EXIT_FAILURE; }
        Vec3i d = volume.getDims(); Vec3f s = volume.getDist(); Vec2f m = volume.getVoxelMapping();
        std::printf("Object: StructuredVolume\n  dims: (%d,%d,%d)\n  dataFormat: %d\n  dist: (%.1f,%.1f,%.1f)\n  voxelMapping: (%.1f,%.1f)\ndata:\n",
                    d.x, d.y, d.z, (int)volume.getDataFormat(), s.x, s.y, s.z, m.x, m.y);
        Vec3i first{ 0, 0, 0 }, last = d;
        if (range > 0) { first = o.first; last = o.last; }
        for (int32_t z = first.z; z < last.z; ++z) {
            std::printf("[%d]\n{\n", z);
            for (int32_t y = first.y; y < last.y; ++y) {
                std::printf("  [%d] {", y);
                for (int32_t x = first.x; x < last.x; ++x) std::printf(x + 1 < last.x ? "%.1f, " : "%.1f", volume.getValue(x, y, z));
                std::printf("}\n");
            }
            std::printf(z + 1 < last.z ? "}\n\n" : "}\n");
        }
    }
    return EXIT_SUCCESS;
}
