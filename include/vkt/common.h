/* C99 API of the volkit-b200 host mirror: shared types (reference include/c/vkt/common.h:9-72, forward.h, hot-path
 * subset).  Numeric enum values equal the C++ API's and the C-ABI's. */
#ifndef VKT_COMMON_H
#define VKT_COMMON_H

#include <stddef.h>
#include <stdint.h>

#ifndef VKTAPI
#if defined(__GNUC__)
#define VKTAPI __attribute__((visibility("default")))
#else
#define VKTAPI
#endif
#endif

typedef uint8_t vktBool_t;
#define VKT_FALSE 0
#define VKT_TRUE  1

typedef enum
{
    vktInvalidValue      = -1,
    vktNoError           =  0,
    vktInvalidDataSource =  1,
    vktReadError         =  2,
    vktWriteError        =  3,
} vktError;

typedef enum
{
    vktDataFormatUnspecified,
    vktDataFormatInt8,
    vktDataFormatInt16,
    vktDataFormatInt32,
    vktDataFormatUInt8,
    vktDataFormatUInt16,
    vktDataFormatUInt32,
    vktDataFormatFloat32,
    vktVoxelFormatCount,
} vktDataFormat;

/* opaque handle (reference include/c/vkt/forward.h) */
struct vktStructuredVolume_impl;
typedef struct vktStructuredVolume_impl* vktStructuredVolume;

#endif
