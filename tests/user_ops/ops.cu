// Example user ops for Transform under the GPU policy: ONE source, compiled by g++ into host functions (run by the
// unmodified reference as the oracle) and by nvcc / NVRTC into sm_100a kernels (tests/test_user_ops.py).
#include <vktb/transform_op.cuh>

// the op of the reference's own example, MakeCheckered<3> (src/examples/Arithmetic.cpp:8-21), restated
VKTB_UNARY_OP(checkered3, x, y, z, voxel)
{
    x >>= 3; y >>= 3; z >>= 3;
    unsigned long long linearIndex = (unsigned long long)(long long)(z * 32 * 32 + y * 32 + x);
    if ((y % 2 == z % 2 && linearIndex % 2 == 0) || (y % 2 != z % 2 && linearIndex % 2 == 1))
        voxel.bytes[0] = 128;
    else
        voxel.bytes[0] = 0;
}

// float arithmetic on a 16-bit voxel: every operation rounds on its own on both sides (--fmad=false / -ffp-contract=off)
VKTB_UNARY_OP(ramp16, x, y, z, voxel)
{
    unsigned v = voxel.bytes[0] | ((unsigned)voxel.bytes[1] << 8);
    float f = (float)v * 0.75f + (float)((x + 2 * y + 3 * z) & 255) * 1.5f;
    f = f < 0.f ? 0.f : (f > 65535.f ? 65535.f : f);
    v = (unsigned)f;
    voxel.bytes[0] = (uint8_t)(v & 0xffu);
    voxel.bytes[1] = (uint8_t)(v >> 8);
}

// binary: the first volume receives the byte-wise maximum, the second the byte-wise minimum (both are modified)
VKTB_BINARY_OP(sort_pair, x, y, z, v1, v2)
{
    const int n = v1.dataFormat == 4 ? 1 : (v1.dataFormat == 5 || v1.dataFormat == 2) ? 2 : 4;   // UInt8 / (U)Int16 / 32-bit
    for (int i = 0; i < n; ++i) {
        const uint8_t a = v1.bytes[i], b = v2.bytes[i];
        v1.bytes[i] = a > b ? a : b;
        v2.bytes[i] = a > b ? b : a;
    }
    (void)x; (void)y; (void)z;
}
