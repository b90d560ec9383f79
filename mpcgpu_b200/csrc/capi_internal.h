// capi_internal.h -- state shared by the translation units of libmpcgpu_b200.so (not exported).
#pragma once
#include <cuda_runtime.h>

namespace mpcb200 {
namespace detail {
__attribute__((visibility("hidden"))) int fail(cudaError_t e, const char* what);   // records the text, returns (int)e
__attribute__((visibility("hidden"))) int bad(const char* what);                   // cudaErrorInvalidValue + text
__attribute__((visibility("hidden"))) void count_launch();
}  // namespace detail
}  // namespace mpcb200
