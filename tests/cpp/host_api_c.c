/* Exercises the C99 host mirror (the .h files under include/vkt), modelled on reference src/examples/CoreAlgorithms.c:
 * Create -> vktFillSV(.1f) -> vktCopyRangeSV -> vktTransformRangeSV1 -> CreateCopy -> vktSafeSumSV.
 * usage: host_api_c <cpu|gpu> <dir>      outputs <dir>/c_out_*.raw */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vkt/Arithmetic.h>
#include <vkt/Copy.h>
#include <vkt/Decompose.h>
#include <vkt/ExecutionPolicy.h>
#include <vkt/Fill.h>
#include <vkt/Resample.h>
#include <vkt/StructuredVolume.h>
#include <vkt/Transform.h>

#define VKT_SAFE_CALL(X) do { if ((X) != vktNoError) { fprintf(stderr, "FAILED: %s\n", #X); return 1; } } while (0)

static void transformOp1(int32_t x, int32_t y, int32_t z, vktVoxelView_t voxel)
{
    voxel.bytes[0] ^= (uint8_t)(x + y + z);
}

static void save(vktStructuredVolume v, const char* dir, const char* name)
{
    char path[1024];
    snprintf(path, sizeof path, "%s/%s", dir, name);
    FILE* f = fopen(path, "wb");
    fwrite(vktStructuredVolumeGetData(v), 1, vktStructuredVolumeGetSizeInBytes(v), f);
    fclose(f);
}

int main(int argc, char** argv)
{
    if (argc < 3) return 2;
    int gpu = strcmp(argv[1], "gpu") == 0;
    vktExecutionPolicy_t cpu = vktGetThreadExecutionPolicy();
    vktExecutionPolicy_t run = cpu;
    run.device = gpu ? vktExecutionPolicyDeviceGPU : vktExecutionPolicyDeviceCPU;

    int32_t dims[3] = { 40, 32, 24 };
    vktStructuredVolume volume, volume2, volume3;
    vktStructuredVolumeCreate(&volume, dims[0], dims[1], dims[2], vktDataFormatUInt8, 1.f, 1.f, 1.f, 0.f, 1.f);
    vktStructuredVolumeCreate(&volume2, dims[0], dims[1], dims[2], vktDataFormatUInt8, 1.f, 1.f, 1.f, 0.f, 1.f);

    vktSetThreadExecutionPolicy(run);
    VKT_SAFE_CALL(vktFillSV(volume, .1f));
    VKT_SAFE_CALL(vktFillSV(volume2, .3f));
    VKT_SAFE_CALL(vktFillRangeSV(volume2, 4, 4, 4, 20, 20, 20, .9f));
    /* copy [10,34)x[4,28)x[2,20) of volume2 to the origin of volume */
    VKT_SAFE_CALL(vktCopyRangeSV(volume, volume2, 10, 4, 2, 34, 28, 20, 0, 0, 0));
    if (gpu) VKT_SAFE_CALL(vktBindDeviceOp1(transformOp1, "xor_coord"));
    VKT_SAFE_CALL(vktTransformRangeSV1(volume, 2, 2, 2, 30, 30, 20, transformOp1));
    vktStructuredVolumeCreateCopy(&volume3, volume);
    /* every volume is a managed resource with a handle of its own */
    if (vktStructuredVolumeGetResourceHandle(volume3) == vktStructuredVolumeGetResourceHandle(volume)
        || vktStructuredVolumeGetResourceHandle(volume) == vktStructuredVolumeGetResourceHandle(volume2)) {
        fprintf(stderr, "two volumes share a resource handle\n");
        return 5;
    }
    VKT_SAFE_CALL(vktSafeSumSV(volume3, volume, volume2));
    VKT_SAFE_CALL(vktAbsDiffRangeSV(volume2, volume, volume3, 1, 1, 1, 39, 31, 23, 0, 0, 0));
    /* BrickDecompose as in reference src/examples/Decompose.c: 16^3 bricks, halo 1 on every side */
    vktArray3D_vktStructuredVolume decomp;
    vktArray3D_vktStructuredVolume_CreateEmpty(&decomp);
    VKT_SAFE_CALL(vktBrickDecomposeResizeSV(decomp, volume3, 16, 16, 16, 1, 1, 1, 1, 1, 1));
    VKT_SAFE_CALL(vktBrickDecomposeSV(decomp, volume3, 16, 16, 16, 1, 1, 1, 1, 1, 1));
    vktSetThreadExecutionPolicy(cpu);

    {
        char path[1024];
        snprintf(path, sizeof path, "%s/%s", argv[2], "c_out_bricks.raw");
        FILE* f = fopen(path, "wb");
        vktVec3i_t nb = vktArray3D_vktStructuredVolume_Dims(decomp);
        if (nb.x != 3 || nb.y != 2 || nb.z != 2 || vktArray3D_vktStructuredVolume_NumElements(decomp) != 12) return 3;
        for (vktStructuredVolume* b = vktArray3D_vktStructuredVolume_Begin(decomp);
             b != vktArray3D_vktStructuredVolume_End(decomp); ++b) {
            fwrite(vktStructuredVolumeGetData(*b), 1, vktStructuredVolumeGetSizeInBytes(*b), f);
            vktStructuredVolumeDestroy(*b);
        }
        fclose(f);
        vktArray3D_vktStructuredVolume_Destroy(decomp);
    }
    save(volume, argv[2], "c_out_volume.raw");
    save(volume2, argv[2], "c_out_volume2.raw");
    save(volume3, argv[2], "c_out_volume3.raw");
    vktStructuredVolumeDestroy(volume);
    vktStructuredVolumeDestroy(volume2);
    vktStructuredVolumeDestroy(volume3);
    printf("ok\n");
    return 0;
}
