// Shared host-side helpers for the C-ABI translation units: error text, CUDA error checks, launch counter.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include "../../include/gigaenv.h"

namespace ggl {

void set_error(const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;

inline void count_launch(int n = 1) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

#define GGL_CUDA(expr)                                                                      \
    do {                                                                                    \
        cudaError_t _e = (expr);                                                            \
        if (_e != cudaSuccess) {                                                            \
            ggl::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return GGL_ERR_CUDA;                                                            \
        }                                                                                   \
    } while (0)

#define GGL_LAUNCH_CHECK()                                                                  \
    do {                                                                                    \
        cudaError_t _e = cudaGetLastError();                                                \
        if (_e != cudaSuccess) {                                                            \
            ggl::set_error("%s:%d kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
            return GGL_ERR_CUDA;                                                            \
        }                                                                                   \
    } while (0)

}  // namespace ggl
