// Shared host-side helpers for libfvp_b200: error reporting and launch checks.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>

#include "../../include/fvp_b200.h"

namespace fvp {

char *error_buffer();  // thread local, 512 bytes (fvp_abi.cu)

inline int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(error_buffer(), 512, fmt, ap);
    va_end(ap);
    return code;
}

inline int check_launch(const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(FVP_E_CUDA, "%s: %s", what, cudaGetErrorString(e));
    return FVP_OK;
}

#define FVP_REQUIRE(cond, ...)                                \
    do {                                                      \
        if (!(cond)) return fvp::fail(FVP_E_ARG, __VA_ARGS__); \
    } while (0)

#define FVP_CUDA(call)                                                                        \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) return fvp::fail(FVP_E_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// -DFVP_PARANOID: device-side bounds assertions on every computed table entry, pixel-record index, occupancy word and
// plane cell (tools/paranoid_check.sh runs the GPU suite against such a build; compute-sanitizer is closed on this pool).
#ifdef FVP_PARANOID
#define FVP_ASSERT(cond)                                                                                   \
    do {                                                                                                   \
        if (!(cond)) {                                                                                     \
            printf("FVP_ASSERT failed: %s  (%s:%d, block %d,%d thread %d)\n", #cond, __FILE__, __LINE__,  \
                   (int)blockIdx.x, (int)blockIdx.y, (int)threadIdx.x);                                    \
            __trap();                                                                                      \
        }                                                                                                  \
    } while (0)
#else
#define FVP_ASSERT(cond) ((void)0)
#endif

}  // namespace fvp
