// Shared helpers for libtmog_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/tmog_b200.h"

namespace tmog {

// thread-local error text behind tmog_last_error()
void set_error(const char* fmt, ...);

#define TMOG_REQUIRE(cond, ...)                       \
    do {                                              \
        if (!(cond)) {                                \
            ::tmog::set_error(__VA_ARGS__);           \
            return TMOG_EINVAL;                       \
        }                                             \
    } while (0)

#define TMOG_CUDA(call)                                                                              \
    do {                                                                                             \
        cudaError_t _e = (call);                                                                     \
        if (_e != cudaSuccess) {                                                                     \
            ::tmog::set_error("CUDA error %s at %s:%d (%s)", cudaGetErrorName(_e), __FILE__, __LINE__, \
                              cudaGetErrorString(_e));                                               \
            return TMOG_ECUDA;                                                                       \
        }                                                                                            \
    } while (0)

#define TMOG_LAUNCH_CHECK() TMOG_CUDA(cudaGetLastError())

int sm_count();  // SMs of the current device (cached per device)

constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(FULL, v, o));
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(FULL, v, o));
    return v;
}

// streaming 128-bit load: read once, keep out of L1
__device__ __forceinline__ double2 ldg_stream(const double2* p) {
    double2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
    return r;
}
// streaming 128-bit store (written once, never re-read by this kernel)
__device__ __forceinline__ void stg_stream(double2* p, double2 v) {
    asm volatile("st.global.cs.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
}

}  // namespace tmog
