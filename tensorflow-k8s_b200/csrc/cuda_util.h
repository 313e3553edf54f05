// cuda_util.h -- CUDA error mapping and per-device facts for the kernel translation units.
#pragma once
#include <cuda_runtime.h>

#include <atomic>

#include "common.h"

#define B200PS_CUDA(expr)                                                                         \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess)                                                                    \
            return ::b200ps::fail(B200PS_ERR_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                                  __FILE__, __LINE__);                                            \
    } while (0)

namespace b200ps {

extern std::atomic<uint64_t> g_launches;  // kernels launched by this library (b200ps_launch_count)

// SM count of the current device (148 on B200), cached per device
inline cudaError_t sm_count(int* out) {
    static int cached[64] = {0};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (cached[dev] == 0) {
        int n = 0;
        e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
        cached[dev] = n;
    }
    *out = cached[dev];
    return cudaSuccess;
}

}  // namespace b200ps
