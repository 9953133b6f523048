// Exercises the C++ host mirror (include/vkt/*.hpp) the way the reference's examples do
// (src/examples/CoreAlgorithms.c, Arithmetic.cpp, ExecutionPolicy.cpp): volumes are created and filled on the CPU,
// the thread's ExecutionPolicy is switched, algorithms run, results are read back after switching to CPU again.
// usage: host_api_cpp <cpu|gpu|gpu-stream|gpu-multi:ids|gpu-oom> <dir>      inputs <dir>/a.raw, b.raw (48x40x36 UInt16); outputs <dir>/out_*.raw
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>

#include <vkt/Aggregates.hpp>
#include <vkt/Arithmetic.hpp>
#include <vkt/Copy.hpp>
#include <vkt/ExecutionPolicy.hpp>
#include <vkt/Fill.hpp>
#include <vkt/Filter.hpp>
#include <vkt/Decompose.hpp>
#include <vkt/Flip.hpp>
#include <vkt/Histogram.hpp>
#include <vkt/InputStream.hpp>
#include <vkt/Memory.hpp>
#include <vkt/OutputStream.hpp>
#include <vkt/RawFile.hpp>
#include <vkt/Resample.hpp>
#include <vkt/Rotate.hpp>
#include <vkt/Scale.hpp>
#include <vkt/Scan.hpp>
#include <vkt/StructuredVolume.hpp>
#include <vkt/Transform.hpp>

using namespace vkt;

static std::string g_dir;

static void load(StructuredVolume& v, const char* name)
{
    FILE* f = std::fopen((g_dir + "/" + name).c_str(), "rb");
    if (!f || std::fread(v.getData(), 1, v.getSizeInBytes(), f) != v.getSizeInBytes()) { std::perror(name); std::exit(2); }
    std::fclose(f);
}

static void save(StructuredVolume& v, const char* name)
{
    FILE* f = std::fopen((g_dir + "/" + name).c_str(), "wb");
    std::fwrite(v.getData(), 1, v.getSizeInBytes(), f);   // getData() under the CPU policy migrates back
    std::fclose(f);
}

#define CHECK(X) do { if ((X) != NoError) { std::fprintf(stderr, "FAILED: %s\n", #X); return 1; } } while (0)

// the ops of reference src/examples/CoreAlgorithms.c:12-24 (unary: first byte ^= x+y+z; binary: v1 |= v2, v2 = v1)
static void xorCoord(int32_t x, int32_t y, int32_t z, VoxelView voxel) { voxel.bytes[0] ^= (uint8_t)(x + y + z); }
static void orAssign(int32_t, int32_t, int32_t, VoxelView v1, VoxelView v2)
{
    int n = v1.dataFormat == DataFormat::UInt8 ? 1 : v1.dataFormat == DataFormat::UInt16 ? 2 : 4;
    for (int i = 0; i < n; ++i) { v1.bytes[i] |= v2.bytes[i]; v2.bytes[i] = v1.bytes[i]; }
}
static void unboundOp(int32_t, int32_t, int32_t, VoxelView) {}

int main(int argc, char** argv)
{
    if (argc < 3) return 2;
    const bool stream = std::strcmp(argv[1], "gpu-stream") == 0;    // GPU policy + host streaming extension
    const bool multi = std::strncmp(argv[1], "gpu-multi", 9) == 0;  // GPU policy + a device group: "gpu-multi:0,0,0"
    const bool gpu = stream || multi || std::strcmp(argv[1], "gpu") == 0;
    g_dir = argv[2];

    ExecutionPolicy cpu = GetThreadExecutionPolicy();            // default: CPU / Serial
    if (cpu.device != ExecutionPolicy::Device::CPU) return 3;
    ExecutionPolicy run = cpu;
    run.device = gpu ? ExecutionPolicy::Device::GPU : ExecutionPolicy::Device::CPU;

    // policies are per thread and not inherited (reference include/cpp/vkt/ExecutionPolicy.hpp:93-100)
    SetThreadExecutionPolicy(run);
    bool childIsDefault = false;
    std::thread([&] { childIsDefault = GetThreadExecutionPolicy().device == ExecutionPolicy::Device::CPU; }).join();
    if (!childIsDefault) return 4;
    SetThreadExecutionPolicy(cpu);

    if (std::strcmp(argv[1], "gpu-oom") == 0) {
        // Fault injection: the allocator fails the migration of `v`; the call must return
        // InvalidValue, the volume must keep its bytes on the host, and the same call must succeed afterwards.
        const int N = 32;
        StructuredVolume v(N, N, N, DataFormat::UInt16), w(N, N, N, DataFormat::UInt16), r(N, N, N, DataFormat::UInt16);
        CHECK(Fill(v, .25f));
        CHECK(Fill(w, .5f));
        ExecutionPolicy g = cpu;
        g.device = ExecutionPolicy::Device::GPU;
        SetThreadExecutionPolicy(g);
        setenv("VKTB_FAIL_ALLOC_AT", "2", 1);                  // r migrates (allocation 1), v cannot (allocation 2)
        Error e = SafeSum(r, v, w);
        unsetenv("VKTB_FAIL_ALLOC_AT");
        if (e != InvalidValue) { std::fprintf(stderr, "expected InvalidValue, got %d\n", (int)e); return 11; }
        if (!v.residesOnHost()) { std::fprintf(stderr, "the volume left the host although its migration failed\n"); return 12; }
        CHECK(SafeSum(r, v, w));                               // no injection: the same call works, nothing was lost
        SetThreadExecutionPolicy(cpu);
        const uint16_t* pr = (const uint16_t*)r.getData();
        const uint16_t* pv = (const uint16_t*)v.getData();
        StructuredVolume want(N, N, N, DataFormat::UInt16);
        CHECK(SafeSum(want, v, w));
        if (std::memcmp(pr, want.getData(), r.getSizeInBytes()) != 0 || pv[0] == 0) return 13;
        std::printf("oom ok\n");
        return 0;
    }

    if (multi) {
        // One process, several GPUs: volumes above the striping threshold (the test lowers it to 1 MiB) are striped over
        // the group at allocation; the unchanged vkt:: calls then run on every device.  256 x 200 x 90 UInt16 = 9.2 MB.
        int ids[16], n = 0;
        for (const char* p = std::strchr(argv[1], ':'); p && *p && n < 16; p = std::strchr(p, ',')) ids[n++] = std::atoi(++p);
        if (n < 2) return 10;
        CHECK(SetDevices(ids, n));
        const int MX = 256, MY = 200, MZ = 90;
        StructuredVolume ma(MX, MY, MZ, DataFormat::UInt16), mb(MX, MY, MZ, DataFormat::UInt16);
        {
            uint16_t* pa = (uint16_t*)ma.getData();
            uint16_t* pb = (uint16_t*)mb.getData();
            for (std::size_t i = 0; i < (std::size_t)MX * MY * MZ; ++i) {
                pa[i] = (uint16_t)((i * 2654435761ull) >> 11);
                pb[i] = (uint16_t)((i * 40503ull + 12345ull) >> 3);
            }
        }
        StructuredVolume msum(MX, MY, MZ, DataFormat::UInt16), mup(384, 300, 135, DataFormat::UInt16);
        SetThreadExecutionPolicy(run);
        CHECK(SafeSum(msum, ma, mb));                          // migrate(): striped H2D on all copy engines, then fan-out
        CHECK(SumRange(msum, ma, mb, Vec3i{ 2, 3, 4 }, Vec3i{ 250, 190, 80 }, Vec3i{ 1, 2, 9 }));
        CHECK(Resample(mup, ma, FilterMode::Linear));
        SetThreadExecutionPolicy(cpu);
        save(msum, "out_m_sum.raw"); save(mup, "out_m_up.raw");
        int one = ids[0];
        (void)one;
    }

    const int X = 48, Y = 40, Z = 36;
    StructuredVolume a(X, Y, Z, DataFormat::UInt16), b(X, Y, Z, DataFormat::UInt16);
    load(a, "a.raw");
    load(b, "b.raw");
    StructuredVolume sum(X, Y, Z, DataFormat::UInt16), diff(X, Y, Z, DataFormat::UInt16);
    StructuredVolume cpy(64, 48, 44, DataFormat::UInt16);
    StructuredVolume prod(X, Y, Z, DataFormat::Float32, 1.f, 1.f, 1.f, -1000.f, 3000.f);

    SetThreadExecutionPolicy(run);      // deferred: nothing moves until an algorithm touches a volume

    if (stream) {
        // host streaming: volumes that all live on the host are streamed through the GPU in z-slabs and STAY on the
        // host; the voxels are those of the migrating call (saved as out_s_*.raw, compared with the same references)
        if (GetHostStreaming()) return 7;                  // off by default
        SetHostStreaming(true);
        SetThreadExecutionPolicy(cpu);                     // volumes are allocated where the policy says: on the host
        StructuredVolume ssum(X, Y, Z, DataFormat::UInt16), sup(72, 60, 54, DataFormat::UInt16);
        StructuredVolume st(a);                            // host copy of a
        SetThreadExecutionPolicy(run);
        CHECK(BindDeviceOp(xorCoord, "xor_coord"));
        CHECK(SafeSum(ssum, a, b));
        CHECK(Resample(sup, a, FilterMode::Linear));
        CHECK(TransformRange(st, 1, 2, 3, 47, 39, 35, xorCoord));
        if (!ssum.residesOnHost() || !sup.residesOnHost() || !st.residesOnHost() || !a.residesOnHost() || !b.residesOnHost()) {
            std::fprintf(stderr, "host streaming migrated a volume\n");
            return 8;
        }
        SetThreadExecutionPolicy(cpu);
        save(ssum, "out_s_sum.raw"); save(sup, "out_s_up.raw"); save(st, "out_s_t.raw");
        SetThreadExecutionPolicy(run);
        // a call with a device-resident volume migrates as usual, streaming on or off
        CHECK(Fill(ssum, .5f));
        if (ssum.residesOnHost()) return 9;
    }

    CHECK(Fill(diff, .25f));
    CHECK(Fill(cpy, .5f));
    CHECK(SafeSum(sum, a, b));
    CHECK(SafeDiffRange(diff, a, b, Vec3i{ 3, 5, 7 }, Vec3i{ 47, 38, 33 }, Vec3i{ 1, 2, 3 }));
    CHECK(Prod(prod, a, b));
    CHECK(CopyRange(cpy, a, Vec3i{ -3, -2, -1 }, Vec3i{ 51, 42, 37 }, Vec3i{ 1, 2, 3 }));   // clamped src

    StructuredVolume t1(a), t2(b);      // copies are made on the current policy's device
    if (gpu) {
        // an unbound host op must fail under the GPU policy instead of running on the CPU
        if (Transform(t1, unboundOp) != InvalidValue) { std::fprintf(stderr, "unbound op did not fail\n"); return 5; }
        CHECK(BindDeviceOp(xorCoord, "xor_coord"));
        CHECK(BindDeviceOp(orAssign, "or_assign"));
    }
    CHECK(TransformRange(t1, 1, 2, 3, 47, 39, 35, xorCoord));
    CHECK(Transform(t1, t2, orAssign));

    StructuredVolume up(72, 60, 54, DataFormat::UInt16), down(24, 20, 18, DataFormat::UInt8, 1.f, 1.f, 1.f, -1.f, 2.f);
    CHECK(Resample(up, a, FilterMode::Linear));
    CHECK(Resample(down, a, FilterMode::Nearest));

    // Flip (as the reference is written: Axis::Y mirrors y and z, Axis::Z mirrors nothing) and Histogram
    StructuredVolume fx(X, Y, Z, DataFormat::UInt16), fy(64, 48, 44, DataFormat::UInt16), fz(a);
    CHECK(Fill(fx, 0.f));
    CHECK(Fill(fy, .5f));
    CHECK(Flip(fx, a, Axis::X));
    CHECK(FlipRange(fy, a, Vec3i{ 2, 1, 3 }, Vec3i{ 46, 39, 35 }, Vec3i{ 5, 4, 3 }, Axis::Y));
    CHECK(Flip(fz, fz, Axis::Z));                       // in place
    Histogram hist(300);
    CHECK(ComputeHistogramRange(a, hist, 1, 2, 3, 47, 39, 35));

    // Scale, Rotate (into pre-filled volumes: voxels that map outside the source keep their value), ApplyFilter
    StructuredVolume scl(X, Y, Z, DataFormat::UInt16), rot(64, 48, 44, DataFormat::UInt16), flt(X, Y, Z, DataFormat::UInt16), fin(a);
    CHECK(Fill(scl, .125f));
    CHECK(Fill(rot, .25f));
    CHECK(Scale(scl, a, Vec3f{ 1.5f, .75f, 2.f }, Vec3f{ 20.f, 18.f, 16.f }));
    CHECK(Rotate(rot, a, Vec3f{ 1.f, 2.f, .5f }, .7f, Vec3f{ 24.f, 20.f, 18.f }));
    float box[27];
    for (int i = 0; i < 27; ++i) box[i] = 1.f / 27.f;
    Filter filter(box, Vec3i{ 3, 3, 3 });
    CHECK(ApplyFilter(flt, a, filter, AddressMode::Clamp));
    CHECK(ApplyFilter(fin, fin, filter, AddressMode::Clamp));        // in place: the reference filters a copy
    float big[125] = {};
    Filter tooBig(big, Vec3i{ 5, 5, 5 });
    if (ApplyFilter(flt, a, tooBig, AddressMode::Clamp) != InvalidValue) return 13;
    if (Scale(fin, fin, Vec3f{ 1.f, 1.f, 1.f }, Vec3f{ 0.f, 0.f, 0.f }) != InvalidValue) return 14;

    // Stream I/O under the running policy: whole volume out and back in, then rows of a range (lines start at x = 0
    // of each row, as the reference addresses them).  RawFile counts items, so the codes are Write/ReadError.
    StructuredVolume ioBack(X, Y, Z, DataFormat::UInt16), ioRange(X, Y, Z, DataFormat::UInt16);
    CHECK(Fill(ioBack, .5f));
    CHECK(Fill(ioRange, .75f));
    {
        std::string whole = g_dir + "/io_48x40x36_uint16.raw", part = g_dir + "/io_part.raw";
        {
            RawFile f(whole.c_str(), "wb");
            if (f.getDims().x != 48 || f.getDims().y != 40 || f.getDims().z != 36 || f.getDataFormat() != DataFormat::UInt16) return 15;
            OutputStream os(f);
            if (os.write(a) != WriteError) return 16;          // the bytes are written; RawFile::write returned 1 item
            CHECK(os.flush());
        }
        {
            RawFile f(whole.c_str(), "rb");
            InputStream is(f);
            if (is.read(ioBack) != ReadError) return 17;       // likewise
        }
        {
            RawFile f(part.c_str(), "wb");
            OutputStream os(f);
            os.writeRange(b, Vec3i{ 3, 1, 2 }, Vec3i{ 41, 38, 30 });
            CHECK(os.flush());
        }
        {
            RawFile f(part.c_str(), "rb");
            InputStream is(f);
            is.readRange(ioRange, Vec3i{ 3, 1, 2 }, Vec3i{ 41, 38, 30 });
        }
    }

    // Aggregates over a range (numElems stays the whole volume's voxel count)
    Aggregates agg;
    CHECK(ComputeAggregatesRange(a, agg, 1, 2, 3, 47, 39, 35));
    if (ComputeAggregatesRange(a, agg, Vec3i{ 5, 2, 3 }, Vec3i{ 1, 39, 35 }) != InvalidValue) return 12;

    // BrickDecompose: 20x16x16 bricks with halos (-1,-2,0)/(2,1,1): 3x3x3 bricks, the far ones cropped
    Array3D<StructuredVolume> bricks;
    CHECK(BrickDecomposeResize(bricks, a, Vec3i{ 20, 16, 16 }, Vec3i{ 1, 2, 0 }, Vec3i{ 2, 1, 1 }));
    if (bricks.dims().x != 3 || bricks.dims().y != 3 || bricks.dims().z != 3) return 10;
    CHECK(BrickDecompose(bricks, a, Vec3i{ 20, 16, 16 }, Vec3i{ 1, 2, 0 }, Vec3i{ 2, 1, 1 }));

    // Scan: out of place with a dst offset, then in place and right to left on x
    StructuredVolume sc(64, 48, 44, DataFormat::UInt16), sci(b);
    CHECK(Fill(sc, .5f));
    CHECK(ScanRange(sc, a, Vec3i{ 2, 1, 3 }, Vec3i{ 46, 39, 35 }, Vec3i{ 5, 4, 3 }));
    CHECK(ScanRange(sci, sci, Vec3i{ 45, 2, 1 }, Vec3i{ 3, 38, 30 }));
    if (ScanRange(sc, a, Vec3i{ 5, 1, 3 }, Vec3i{ 2, 39, 35 }) != InvalidValue) return 11;

    // single-voxel accessors work under either policy
    sum.setValue(1, 2, 3, .75f);
    float back = sum.getValue(1, 2, 3);
    if (back < .7499f || back > .7501f) { std::fprintf(stderr, "setValue/getValue round trip: %f\n", back); return 6; }

    SetThreadExecutionPolicy(cpu);
    save(sum, "out_sum.raw"); save(diff, "out_diff.raw"); save(prod, "out_prod.raw"); save(cpy, "out_cpy.raw");
    save(t1, "out_t1.raw"); save(t2, "out_t2.raw"); save(up, "out_up.raw"); save(down, "out_down.raw");
    save(sc, "out_sc.raw"); save(sci, "out_sci.raw");
    save(ioBack, "out_ioback.raw"); save(ioRange, "out_iorange.raw");
    save(scl, "out_scl.raw"); save(rot, "out_rot.raw"); save(flt, "out_flt.raw"); save(fin, "out_fin.raw");
    save(fx, "out_fx.raw"); save(fy, "out_fy.raw"); save(fz, "out_fz.raw");
    {
        FILE* f = std::fopen((g_dir + "/out_hist.raw").c_str(), "wb");
        std::fwrite(hist.getBinCounts(), sizeof(std::size_t), hist.getNumBins(), f);   // migrates the bins back
        std::fclose(f);
    }
    {
        FILE* f = std::fopen((g_dir + "/out_agg.raw").c_str(), "wb");
        std::fwrite(&agg, sizeof(agg), 1, f);
        std::fclose(f);
    }
    {
        FILE* f = std::fopen((g_dir + "/out_bricks.raw").c_str(), "wb");
        for (StructuredVolume& b : bricks) std::fwrite(b.getData(), 1, b.getSizeInBytes(), f);
        std::fclose(f);
    }
    std::printf("ok\n");
    return 0;
}
