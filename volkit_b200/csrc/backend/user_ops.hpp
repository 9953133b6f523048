// User-supplied Transform ops: handles >= kUserOpBase name kernels loaded from a user image (user_ops.cu).
#pragma once

#include <cuda_runtime.h>

#include "../kernels/launch.hpp"

namespace vktb {

constexpr int kUserOpBase = 1000;

// text of include/vktb/transform_op.cuh, embedded at build time (build/transform_op_text.inc) for NVRTC
extern const char* const kTransformOpHeaderText;

bool userOpValid(int op, bool binary);
// v2 == nullptr for a unary op
cudaError_t launchUserTransform(const LaunchCtx&, int op, const VolDesc& v1, const VolDesc* v2, const int32_t first[3],
                                const int32_t last[3]);

} // namespace vktb
