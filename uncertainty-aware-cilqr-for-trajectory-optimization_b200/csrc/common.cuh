// common.cuh — error plumbing shared by the two translation units of libcilqr_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include "../../include/cilqr.h"

namespace cilqr
{
extern thread_local char g_err[512];
extern std::atomic<long long> g_launches;

inline int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define CILQR_CUDA(expr)                                                                                      \
    do {                                                                                                      \
        cudaError_t e__ = (expr);                                                                             \
        if (e__ != cudaSuccess)                                                                               \
            return ::cilqr::fail(CILQR_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

// Every kernel launch of the library goes through this so that bench.py can report gpu_launches.
#define CILQR_LAUNCH_CHECK()                                                                                  \
    do {                                                                                                      \
        ::cilqr::g_launches.fetch_add(1, std::memory_order_relaxed);                                          \
        cudaError_t e__ = cudaGetLastError();                                                                 \
        if (e__ != cudaSuccess)                                                                               \
            return ::cilqr::fail(CILQR_E_CUDA, "kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

#ifdef __CUDACC__
// GridMapRosConverter::toOccupancyGrid, one cell (grid_map_ros/src/GridMapRosConverter.cpp:297-302):
// value = isnan(v) ? -1 : 0 + min(max(0, v), 1) * 100 with v = (data - min) / (max - min), assigned to an int8 (truncation).
__device__ __forceinline__ int8_t occupancy_cell(float d, float dmin, float dmax)
{
    float value = __fdiv_rn(__fsub_rn(d, dmin), __fsub_rn(dmax, dmin));
    if (isnan(value))
        value = -1.0f;
    else
        value = __fadd_rn(0.0f, __fmul_rn(fminf(fmaxf(0.0f, value), 1.0f), 100.0f));
    return (int8_t)(int)value; // float -> integer conversion truncates, like the reference's assignment to int8_t
}
#endif

// Device buffer that is only ever grown (workspaces live as long as the handle).
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes)
    {
        if (bytes <= cap)
            return 0;
        if (p)
            cudaFree(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess)
            return fail(CILQR_E_ALLOC, "cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
        cap = bytes;
        return 0;
    }
    void release()
    {
        if (p)
            cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T>
    T *as() const { return static_cast<T *>(p); }
};
} // namespace cilqr
