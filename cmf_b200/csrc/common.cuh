// Shared helpers for the cmf_b200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>

#include "../../include/cmf_b200.h"

namespace cmf {

void set_error(const char* fmt, ...);
void count_launch();

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

#define CMF_REQUIRE(cond, ...)            \
    do {                                  \
        if (!(cond)) {                    \
            cmf::set_error(__VA_ARGS__);  \
            return -1;                    \
        }                                 \
    } while (0)

#define CMF_CHECK_LAUNCH()                                                  \
    do {                                                                    \
        cmf::count_launch();                                                \
        cudaError_t e__ = cudaPeekAtLastError();                            \
        if (e__ != cudaSuccess) {                                           \
            cmf::set_error("%s:%d: %s", __FILE__, __LINE__, cudaGetErrorString(e__)); \
            return (int)e__;                                                \
        }                                                                   \
    } while (0)

__host__ __device__ constexpr long long ceil_div(long long a, long long b) { return (a + b - 1) / b; }

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
// streaming (evict-first) 128-bit accesses for data touched exactly once
__device__ __forceinline__ float4 ldcs4(const float* p) { return __ldcs(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void stcs4(float* p, float4 v) { __stcs(reinterpret_cast<float4*>(p), v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// "already configured on this device" flag: function attributes (the opt-in to > 48 KB of dynamic
// shared memory) are per device, so one process launching on a second GPU must set them there too.
struct DeviceOnce {
    std::atomic<unsigned long long> mask{0};
    static int current() { int d = -1; cudaGetDevice(&d); return d; }
    bool done() const {
        const int d = current();
        return d >= 0 && d < 64 && ((mask.load(std::memory_order_acquire) >> d) & 1ull);
    }
    void mark() {
        const int d = current();
        if (d >= 0 && d < 64) mask.fetch_or(1ull << d, std::memory_order_release);
    }
};

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace cmf
