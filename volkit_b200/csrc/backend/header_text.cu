// include/vktb/transform_op.cuh as a string constant (generated: build/transform_op_text.inc, see the Makefile).
#include "user_ops.hpp"

namespace vktb {
const char* const kTransformOpHeaderText =
#include "transform_op_text.inc"
    ;
} // namespace vktb
