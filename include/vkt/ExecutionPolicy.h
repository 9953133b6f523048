/* Per-thread execution policy, C API (reference include/c/vkt/ExecutionPolicy.h:13-45). */
#ifndef VKT_EXECUTIONPOLICY_H
#define VKT_EXECUTIONPOLICY_H

#include "common.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef enum { vktExecutionPolicyDeviceCPU, vktExecutionPolicyDeviceGPU } vktExecutionPolicyDevice;
typedef enum { vktExecutionPolicyHostAPISerial, vktExecutionPolicyAPIOpenMP, vktExecutionPolicyAPIAuto } vktExecutionPolicyHostAPI;
typedef enum { vktExecutionPolicyDeviceAPICUDA, vktExecutionPolicyDeviceAPIAuto } vktExecutionPolicyDeviceAPI;

typedef struct
{
    vktExecutionPolicyDevice device;
    vktExecutionPolicyHostAPI hostApi;
    vktExecutionPolicyDeviceAPI deviceApi;
    vktBool_t printPerformance;
} vktExecutionPolicy_t;

VKTAPI void vktSetThreadExecutionPolicy(vktExecutionPolicy_t policy);
VKTAPI vktExecutionPolicy_t vktGetThreadExecutionPolicy(void);

/* volkit-b200 extension: stream host-resident volumes through the GPU in z-slabs instead of migrating them
 * (see vkt::SetHostStreaming in ExecutionPolicy.hpp) */
VKTAPI void vktSetHostStreaming(vktBool_t enabled);
VKTAPI vktBool_t vktGetHostStreaming(void);

#ifdef __cplusplus
}
#endif
#endif
