// Shared helpers for the pinot_b200 C-ABI library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

namespace pb {

// Thread-local last-error string, returned by pinot_b200_last_error().
void set_error(const char* fmt, ...);

// Error codes returned by every extern "C" entry point.
enum : int {
    PB_OK = 0,
    PB_ERR_ARG = 1,      // bad argument (null pointer, negative size, unsupported shape)
    PB_ERR_CUDA = 2,     // CUDA runtime error on launch
    PB_ERR_LIMIT = 3,    // size exceeds a documented limit
};

// Number of kernels this library has launched in the process (bench.py's gpu_launches claim).
void count_launch();

inline int check_launch(const char* what) {
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
        return PB_ERR_CUDA;
    }
    return PB_OK;
}

#define PB_REQUIRE(cond, ...)                \
    do {                                     \
        if (!(cond)) {                       \
            pb::set_error(__VA_ARGS__);      \
            return pb::PB_ERR_ARG;           \
        }                                    \
    } while (0)

#define PB_CUDA(call)                                                                     \
    do {                                                                                  \
        cudaError_t e__ = (call);                                                         \
        if (e__ != cudaSuccess) {                                                         \
            pb::set_error("%s: CUDA error %d (%s)", #call, (int)e__, cudaGetErrorString(e__)); \
            return pb::PB_ERR_CUDA;                                                       \
        }                                                                                 \
    } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

__device__ __forceinline__ float4 ldg4(const float* p) {
    return __ldg(reinterpret_cast<const float4*>(p));
}

}  // namespace pb
