// gpuassert.cuh -- drop-in for the header of that name in the (absent) GBD-PCG submodule.
// gpuErrchk(x): print + abort on any CUDA error.  The reference uses it both with and without
// a trailing ';' (include/qdldl/sqp.cuh:269 has none), so the body is a braced block.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

inline void gpuAssert(cudaError_t code, const char* file, int line, bool abort_on_error = true) {
    if (code != cudaSuccess) {
        fprintf(stderr, "GPUassert: %s %s %d\n", cudaGetErrorString(code), file, line);
        if (abort_on_error) exit((int)code);
    }
}
#define gpuErrchk(ans) { gpuAssert((ans), __FILE__, __LINE__); }
