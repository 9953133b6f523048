// User-supplied Transform ops (include/vktb/transform_op.cuh): registry of device images and their launch.
//
// The reference's contract is an arbitrary TransformUnaryOp / TransformBinaryOp (include/cpp/vkt/Transform.hpp:16-25).
// The user's op, written once with VKTB_UNARY_OP / VKTB_BINARY_OP, compiles for the device into a kernel
// `vktb_unary_<name>` / `vktb_binary_<name>` that carries the op inlined.  An image (cubin, PTX or fatbin) is loaded
// with the runtime's library API (cudaLibraryLoadData: context-independent, so one registration serves every device);
// CUDA C source goes through NVRTC first when libnvrtc can be dlopen'ed (it is optional: the library does not link it).
#include "context.hpp"

#include <dlfcn.h>

#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../../include/vktb/backend.h"
#include "../../../include/vktb/transform_op.cuh"
#include "../kernels/launch.hpp"
#include "user_ops.hpp"

namespace vktb {

namespace {

struct UserOp
{
    cudaLibrary_t lib;
    cudaKernel_t  kernel;
    std::string   name;
    bool          binary;
};

std::mutex g_userMutex;
std::vector<UserOp> g_userOps;

int registerImage(const void* image, size_t size, const char* name, bool binary, int* op)
{
    const char* fn = binary ? "vktb_register_binary_op_image" : "vktb_register_unary_op_image";
    if (!image || !size || !name || !op) return fail((std::string(fn) + ": null argument").c_str());
    if (!currentDevice()) return -1;
    // PTX text must be NUL-terminated for the loader: copy every image into a terminated buffer
    std::vector<char> buf((const char*)image, (const char*)image + size);
    buf.push_back('\0');
    cudaLibrary_t lib = nullptr;
    cudaError_t e = cudaLibraryLoadData(&lib, buf.data(), nullptr, nullptr, 0, nullptr, nullptr, 0);
    if (e != cudaSuccess) return fail((std::string(fn) + ": cudaLibraryLoadData (is the image built for sm_100a?)").c_str(), e);
    const std::string kname = std::string(binary ? "vktb_binary_" : "vktb_unary_") + name;
    cudaKernel_t k = nullptr;
    e = cudaLibraryGetKernel(&k, lib, kname.c_str());
    if (e != cudaSuccess) {
        cudaLibraryUnload(lib);
        return fail((std::string(fn) + ": the image has no kernel '" + kname + "' (VKTB_" + (binary ? "BINARY" : "UNARY")
                     + "_OP(" + name + ", ...) of vktb/transform_op.cuh defines it)").c_str(), e);
    }
    std::lock_guard<std::mutex> lock(g_userMutex);
    g_userOps.push_back({ lib, k, name, binary });
    *op = kUserOpBase + (int)g_userOps.size() - 1;
    return 0;
}

// ---- NVRTC, loaded on demand --------------------------------------------------------------------------------------
struct Nvrtc
{
    void* so = nullptr;
    int (*createProgram)(void**, const char*, const char*, int, const char* const*, const char* const*) = nullptr;
    int (*compileProgram)(void*, int, const char* const*) = nullptr;
    int (*getCUBINSize)(void*, size_t*) = nullptr;
    int (*getCUBIN)(void*, char*) = nullptr;
    int (*getLogSize)(void*, size_t*) = nullptr;
    int (*getLog)(void*, char*) = nullptr;
    int (*destroyProgram)(void**) = nullptr;
    bool ok = false;
};

Nvrtc& nvrtc()
{
    static Nvrtc n;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const char* name : { "libnvrtc.so.12", "libnvrtc.so", "libnvrtc.so.13" }) {
            n.so = dlopen(name, RTLD_NOW | RTLD_LOCAL);
            if (n.so) break;
        }
        if (!n.so) return;
        auto sym = [&](const char* s) { return dlsym(n.so, s); };
        n.createProgram = (decltype(n.createProgram))sym("nvrtcCreateProgram");
        n.compileProgram = (decltype(n.compileProgram))sym("nvrtcCompileProgram");
        n.getCUBINSize = (decltype(n.getCUBINSize))sym("nvrtcGetCUBINSize");
        n.getCUBIN = (decltype(n.getCUBIN))sym("nvrtcGetCUBIN");
        n.getLogSize = (decltype(n.getLogSize))sym("nvrtcGetProgramLogSize");
        n.getLog = (decltype(n.getLog))sym("nvrtcGetProgramLog");
        n.destroyProgram = (decltype(n.destroyProgram))sym("nvrtcDestroyProgram");
        n.ok = n.createProgram && n.compileProgram && n.getCUBINSize && n.getCUBIN && n.getLogSize && n.getLog && n.destroyProgram;
    });
    return n;
}

int registerSource(const char* source, const char* name, bool binary, int* op)
{
    const char* fn = binary ? "vktb_register_binary_op_source" : "vktb_register_unary_op_source";
    if (!source || !name || !op) return fail((std::string(fn) + ": null argument").c_str());
    Nvrtc& n = nvrtc();
    if (!n.ok) return fail((std::string(fn) + ": libnvrtc is not available on this machine; compile the op with nvcc "
                            "-cubin -gencode arch=compute_100a,code=sm_100a and use the _image entry point").c_str());
    // the op source includes <vktb/transform_op.cuh>: handed to NVRTC as an in-memory header
    const char* headers[] = { kTransformOpHeaderText };
    const char* includeNames[] = { "vktb/transform_op.cuh" };
    void* prog = nullptr;
    if (n.createProgram(&prog, source, "vktb_user_op.cu", 1, headers, includeNames) != 0)
        return fail((std::string(fn) + ": nvrtcCreateProgram failed").c_str());
    // --fmad=false: the reference's CPU path has no FMA; a float op must round like the host twin
    const char* opts[] = { "--gpu-architecture=sm_100a", "--fmad=false", "--std=c++17", "-default-device" };
    const int rc = n.compileProgram(prog, 3, opts);
    if (rc != 0) {
        size_t ls = 0;
        n.getLogSize(prog, &ls);
        std::string log(ls ? ls : 1, '\0');
        if (ls) n.getLog(prog, &log[0]);
        n.destroyProgram(&prog);
        return fail((std::string(fn) + ": NVRTC compilation failed:\n" + log).c_str());
    }
    size_t cs = 0;
    n.getCUBINSize(prog, &cs);
    std::vector<char> cubin(cs);
    n.getCUBIN(prog, cubin.data());
    n.destroyProgram(&prog);
    return registerImage(cubin.data(), cubin.size(), name, binary, op);
}

} // namespace

bool userOpValid(int op, bool binary)
{
    std::lock_guard<std::mutex> lock(g_userMutex);
    const int i = op - kUserOpBase;
    return i >= 0 && i < (int)g_userOps.size() && g_userOps[(size_t)i].binary == binary;
}

static vktb_transform_vol describeOperand(const VolDesc& v, const int32_t first[3])
{
    const Operand o = operandAt(v, first[0], first[1], first[2]);
    vktb_transform_vol t;
    t.base = o.base; t.pitchY = o.pitchY; t.pitchZ = o.pitchZ;
    t.bpv = bytesPerVoxel(v.format); t.format = v.format; t.lo = v.map.lo; t.hi = v.map.hi;
    return t;
}

cudaError_t launchUserTransform(const LaunchCtx& c, int op, const VolDesc& v1, const VolDesc* v2, const int32_t first[3],
                                const int32_t last[3])
{
    cudaKernel_t kernel;
    bool binary;
    {
        std::lock_guard<std::mutex> lock(g_userMutex);
        const int i = op - kUserOpBase;
        if (i < 0 || i >= (int)g_userOps.size()) return cudaErrorInvalidValue;
        kernel = g_userOps[(size_t)i].kernel;
        binary = g_userOps[(size_t)i].binary;
    }
    if (binary != (v2 != nullptr)) return cudaErrorInvalidValue;
    vktb_transform_params p;
    std::memset(&p, 0, sizeof p);
    p.nx = last[0] - first[0]; p.ny = last[1] - first[1]; p.nz = last[2] - first[2];
    if (p.nx <= 0 || p.ny <= 0 || p.nz <= 0) return cudaSuccess;
    p.x0 = first[0]; p.y0 = first[1]; p.z0 = first[2];
    p.v1 = describeOperand(v1, first);
    if (v2) p.v2 = describeOperand(*v2, first);
    const long long rows = (long long)p.ny * p.nz;
    const int perThread = binary ? 1 : 16 / p.v1.bpv;
    // rows that all start on a 16-byte boundary need exactly ceil(nx / V) chunks; otherwise one more for the head
    // (an extra, almost empty column of CTAs halves the throughput when a row is exactly one CTA wide)
    const bool alignedRows = ((reinterpret_cast<uintptr_t>(p.v1.base) | (uintptr_t)p.v1.pitchY | (uintptr_t)p.v1.pitchZ) & 15u) == 0;
    const long long items = binary ? p.nx : (p.nx + perThread - 1) / perThread + (alignedRows ? 0 : 1);
    dim3 block(256, 1, 1);
    dim3 grid((unsigned)((items + 255) / 256), (unsigned)(rows > 65535 ? 65535 : rows), 1);
    void* args[] = { &p };
    cudaError_t e = cudaLaunchKernel((const void*)kernel, grid, block, args, 0, c.stream);
    countLaunch();
    return e;
}

} // namespace vktb

using namespace vktb;

extern "C" {

int vktb_register_unary_op_image(const void* image, size_t size, const char* name, int* op)
{
    VKTB_API_RANGE();
    return registerImage(image, size, name, false, op);
}
int vktb_register_binary_op_image(const void* image, size_t size, const char* name, int* op)
{
    VKTB_API_RANGE();
    return registerImage(image, size, name, true, op);
}
int vktb_register_unary_op_source(const char* source, const char* name, int* op) { return registerSource(source, name, false, op); }
int vktb_register_binary_op_source(const char* source, const char* name, int* op) { return registerSource(source, name, true, op); }
int vktb_nvrtc_available(void) { return nvrtc().ok ? 1 : 0; }

} // extern "C"
